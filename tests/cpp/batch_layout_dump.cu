// Host-side dump of the batched kernel's shared-memory layouts (csrc/ekf_batch.cuh) for tests/test_batch_layout.py:
// "n dd" on the command line -> one line per stored element "i c offset", then "S doubles", "SMEM bytes(m=20)".
#include <cstdio>
#include <cstdlib>

#include "ekf_batch.cuh"

int main(int argc, char **argv) {
  const int n = argc > 1 ? atoi(argv[1]) : 43, dd = argc > 2 ? atoi(argv[2]) : 2;
  for (int i = 0; i < n; ++i)
    for (int c = 0; c <= i; ++c)
      printf("%d %d %d\n", i, c, dd == 2 ? ekf::batch_s_idx<2>(i, c) : ekf::batch_s_idx<0>(i, c));
  printf("S %d\n", ekf::batch_s_doubles(n, dd));
  printf("SMEM %zu\n", ekf::batch_smem_bytes(n, 20, dd));
  for (int I = 0; I < 2; ++I)
    for (int l = 0; l < 32; ++l) printf("CH %d %d %d\n", I, l, ekf::batch_tile_chunk(I, l >> 2, l & 3));
  return 0;
}
