// Host harness exposing csrc/dc_core.h to Python (ctypes) for CPU unit tests.
#include "../distributed-kfac-pytorch_b200/csrc/dc_core.h"

extern "C" {
int dc_deflate(int m, double rho, double* ds, double* zs, int* keep, int* defl, int* rot_i, int* rot_j, double* rot_c,
               double* rot_s, int* counts) {
  kfac::DeflateCounts c = kfac::deflate_scan(m, rho, ds, zs, keep, defl, rot_i, rot_j, rot_c, rot_s);
  counts[0] = c.k;
  counts[1] = c.ndefl;
  counts[2] = c.nrot;
  return 0;
}
int dc_secular(const double* dk, const double* z2, double rho, int k, int* origin, double* mu) {
  for (int j = 0; j < k; ++j) kfac::secular_root(dk, z2, rho, k, j, origin[j], mu[j]);
  return 0;
}
}
