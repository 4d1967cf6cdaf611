// Host construction of the L2 sweep order (fc_sweep_order.hpp).
#include "fc_sweep_order.hpp"

#include <algorithm>
#include <cmath>
#include <cstdint>

namespace fc_host
{
  bool l2_sweep_order(int n_col_nodes, const double* X, int n_ele, int nen, const int* conn, long long layer_elems, std::vector<int>& perm)
  {
    perm.clear();
    if (n_ele <= 0 || layer_elems <= 0 || n_ele <= 2 * layer_elems) return false;
    double lo[3] = {1e300, 1e300, 1e300}, hi[3] = {-1e300, -1e300, -1e300}, hsum[3] = {0, 0, 0};
    for (int n = 0; n < n_col_nodes; ++n)
      for (int d = 0; d < 3; ++d)
      {
        lo[d] = std::min(lo[d], X[3 * (size_t)n + d]);
        hi[d] = std::max(hi[d], X[3 * (size_t)n + d]);
      }
    for (int e = 0; e < n_ele; ++e)
      for (int d = 0; d < 3; ++d)
      {
        double mn = 1e300, mx = -1e300;
        for (int a = 0; a < nen; ++a)
        {
          const double x = X[3 * (size_t)conn[(size_t)e * nen + a] + d];
          mn = std::min(mn, x);
          mx = std::max(mx, x);
        }
        hsum[d] += mx - mn;
      }
    long long nd[3];
    double hd[3];
    for (int d = 0; d < 3; ++d)
    {
      const double ext = hi[d] - lo[d], h = hsum[d] / n_ele;
      nd[d] = (ext > 0.0 && h > 0.0) ? std::max(1LL, std::llround(ext / h)) : 1;
      nd[d] = std::min(nd[d], 1LL << 20);
      hd[d] = ext > 0.0 ? ext / (double)nd[d] : 1.0;
    }
    // elements per lattice cell (6 for a Kuhn-split tet mesh)
    const long long per_cell = std::max(1LL, (long long)std::llround((double)n_ele / ((double)nd[0] * (double)nd[1] * (double)nd[2])));
    if (nd[0] * nd[1] * per_cell <= layer_elems) return false;  // a z layer fits: keep the caller's order
    const long long W = std::max(4LL, layer_elems / (nd[0] * per_cell));
    if (W >= nd[1]) return false;
    std::vector<std::pair<unsigned long long, int>> keyed((size_t)n_ele);
    for (int e = 0; e < n_ele; ++e)
    {
      long long q[3];
      for (int d = 0; d < 3; ++d)
      {
        double c = 0.0;
        for (int a = 0; a < nen; ++a) c += X[3 * (size_t)conn[(size_t)e * nen + a] + d];
        q[d] = (long long)std::floor((c / nen - lo[d]) / hd[d]);
        q[d] = std::max(0LL, std::min(nd[d] - 1, q[d]));
      }
      const unsigned long long key = (((unsigned long long)(q[1] / W) * (unsigned long long)nd[2] + (unsigned long long)q[2]) * (unsigned long long)nd[1] +
                                         (unsigned long long)q[1]) * (unsigned long long)nd[0] + (unsigned long long)q[0];
      keyed[e] = {key, e};
    }
    std::sort(keyed.begin(), keyed.end());
    perm.resize((size_t)n_ele);
    bool identity = true;
    for (int k = 0; k < n_ele; ++k)
    {
      perm[k] = keyed[k].second;
      identity = identity && perm[k] == k;
    }
    if (identity) perm.clear();
    return !identity;
  }
}  // namespace fc_host

// ---- C ABI (include/fourc_b200.h) -----------------------------------------------------------------------------
extern "C"
{
  int fc_l2_sweep_order(int n_col_nodes, const double* X, int n_ele, int nen, const int* conn, int64_t layer_elems, int* perm)
  {
    if (!X || !conn || !perm || n_ele < 0) return -1;
    std::vector<int> p;
    const bool changed = fc_host::l2_sweep_order(n_col_nodes, X, n_ele, nen, conn, layer_elems, p);
    for (int k = 0; k < n_ele; ++k) perm[k] = changed ? p[k] : k;
    return changed ? 1 : 0;
  }
}
