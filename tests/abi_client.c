/* A plain C99 client of include/b200dlrm.h: what a cgo / JNI / FFI binding sees.  Built and run by
 * tests/test_abi.py on the CPU container: the header must compile as C (-std=c99 -pedantic), the library must
 * link without torch or Python, and the host-side argument checks must answer before any CUDA call. */
#include <stdio.h>
#include <string.h>

#include "b200dlrm.h"

int main(void) {
  int fails = 0;
  if (b200_version() < 100) ++fails;
  /* negative size -> B200 "invalid argument" status with a message */
  if (b200_emb_fwd(0, 0, 0, 0, 0, 0, -1, 4, 8, 0, 0, B200_F32, B200_F32, 0, 0) != 1) ++fails;
  if (strstr(b200_last_error(), "negative") == 0) ++fails;
  /* workspace sizes are pure host arithmetic */
  if (b200_emb_bwd_workspace_bytes(1000, 128) < 6 * 4000) ++fails;
  if (b200_sort_workspace_bytes(1000) == 0) ++fails;
  if (b200_relu_bwd_bias_workspace_bytes(32768, 1024) < 256u * 1024u * 4u) ++fails;
  /* record wider than the staging buffer */
  if (b200_criteo_transform(0, 8, 100, 13, 0, 0, 26, 0, 8, 0, 13, 0, 0, 0, 0) != 1) ++fails;
  /* the no-alloc sort rejects oversize requests without launching anything */
  if (b200_sort_dedup_noalloc(0, 0, (int64_t)B200_SORT_NOALLOC_MAX_KEYS + 1, 1u << 20, 0, 0, 0, 0, 0, 0) == 0) ++fails;
  printf("abi_client: %d failure(s)\n", fails);
  return fails;
}
