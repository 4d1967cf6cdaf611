// Host construction of the brick plan (fc_brick_plan.hpp).  Pure integer code after the lattice detection.
#include "fc_brick_plan.hpp"

#include <algorithm>
#include <cmath>
#include <numeric>

namespace fc_host
{
  bool lattice_axis(int n_nodes, const double* X, int d, int n_ele, const int* conn, std::vector<int>& idx, int& n_levels);
}

namespace
{
  using fc_host::BrickPlanHost;
  using fc_host::lattice_axis;
}  // namespace

namespace fc_host
{
  // lattice index of every node along axis d: uniform spacing first (O(n)), sorted distinct levels otherwise
  bool lattice_axis(int n_nodes, const double* X, int d, int n_ele, const int* conn, std::vector<int>& idx, int& n_levels)
  {
    double lo = 1e300, hi = -1e300;
    for (int n = 0; n < n_nodes; ++n)
    {
      lo = std::min(lo, X[3 * (size_t)n + d]);
      hi = std::max(hi, X[3 * (size_t)n + d]);
    }
    const double ext = hi - lo;
    idx.assign((size_t)n_nodes, 0);
    if (!(ext > 0.0)) return false;  // flat in this direction: not a mesh of hex cells
    double hsum = 0.0;
    for (int e = 0; e < n_ele; ++e)
    {
      double mn = 1e300, mx = -1e300;
      for (int a = 0; a < 8; ++a)
      {
        const double x = X[3 * (size_t)conn[(size_t)e * 8 + a] + d];
        mn = std::min(mn, x);
        mx = std::max(mx, x);
      }
      hsum += mx - mn;
    }
    const double h0 = hsum / std::max(1, n_ele);
    if (h0 > 0.0 && ext / h0 < 4e6)
    {
      const long long nd = std::max(1LL, std::llround(ext / h0));
      const double h = ext / (double)nd;
      bool ok = true;
      for (int n = 0; n < n_nodes && ok; ++n)
      {
        const double r = (X[3 * (size_t)n + d] - lo) / h;
        const long long k = std::llround(r);
        ok = std::fabs(r - (double)k) <= 1e-6 && k >= 0 && k <= nd;
        idx[n] = (int)k;
      }
      if (ok)
      {
        n_levels = (int)nd + 1;
        return true;
      }
    }
    // non-uniform rectilinear grid: distinct coordinate levels
    std::vector<double> v((size_t)n_nodes);
    for (int n = 0; n < n_nodes; ++n) v[n] = X[3 * (size_t)n + d];
    std::sort(v.begin(), v.end());
    const double tol = 1e-9 * ext;
    std::vector<double> lev;
    for (double x : v)
      if (lev.empty() || x - lev.back() > tol) lev.push_back(x);
    if (lev.size() > (size_t)1 << 22) return false;
    for (int n = 0; n < n_nodes; ++n)
    {
      const double x = X[3 * (size_t)n + d];
      size_t k = (size_t)(std::lower_bound(lev.begin(), lev.end(), x - tol) - lev.begin());
      if (k >= lev.size() || std::fabs(lev[k] - x) > tol) return false;
      idx[n] = (int)k;
    }
    n_levels = (int)lev.size();
    return true;
  }

  bool build_brick_plan_host(int n_col_nodes, const double* X, int n_ele, const int* conn, int n_row_nodes, int max_row_nodes,
      int chunk, BrickPlanHost& plan)
  {
    plan = BrickPlanHost();
    constexpr int BY = BrickPlanHost::BY, BZ = BrickPlanHost::BZ, NL = BrickPlanHost::NL, CY = BrickPlanHost::CY, NE = BrickPlanHost::NE;
    if (n_ele <= 0 || n_row_nodes <= 0 || max_row_nodes > 27) return false;
    // ---- lattice coordinates ------------------------------------------------------------------------------------
    std::vector<int> L[3];
    int nl[3];
    for (int d = 0; d < 3; ++d)
      if (!lattice_axis(n_col_nodes, X, d, n_ele, conn, L[d], nl[d])) return false;
    // sweep axis = the axis along which consecutive owned nodes are lattice neighbours (rows adjacent in memory)
    int ax[3] = {0, 1, 2};
    if (n_row_nodes > 1)
    {
      int diff = -1, cnt = 0;
      for (int d = 0; d < 3; ++d)
        if (L[d][0] != L[d][1])
        {
          diff = d;
          ++cnt;
        }
      if (cnt == 1 && diff != 0) std::swap(ax[0], ax[diff]);
    }
    const std::vector<int>&Lx = L[ax[0]], &Ly = L[ax[1]], &Lz = L[ax[2]];
    const long long nx = nl[ax[0]], ny = nl[ax[1]], nz = nl[ax[2]];
    if (nx * ny * nz > 8LL * n_col_nodes + 4096) return false;  // sparse occupation: not a box-like mesh
    std::vector<int> lat2node((size_t)(nx * ny * nz), -1);
    auto lat = [&](long long i, long long j, long long k) -> size_t { return (size_t)((k * ny + j) * nx + i); };
    for (int n = 0; n < n_col_nodes; ++n)
    {
      int& slot = lat2node[lat(Lx[n], Ly[n], Lz[n])];
      if (slot >= 0) return false;  // two nodes on one lattice point
      slot = n;
    }
    // ---- every element spans exactly one lattice cell ------------------------------------------------------------
    const long long cx = nx - 1, cy = ny - 1, cz = nz - 1;
    if (cx < 1 || cy < 1 || cz < 1) return false;
    std::vector<int> cell2ele((size_t)(cx * cy * cz), -1);
    auto cell = [&](long long i, long long j, long long k) -> size_t { return (size_t)((k * cy + j) * cx + i); };
    for (int e = 0; e < n_ele; ++e)
    {
      const int* en = conn + (size_t)e * 8;
      int m[3] = {1 << 30, 1 << 30, 1 << 30};
      for (int a = 0; a < 8; ++a)
      {
        m[0] = std::min(m[0], Lx[en[a]]);
        m[1] = std::min(m[1], Ly[en[a]]);
        m[2] = std::min(m[2], Lz[en[a]]);
      }
      unsigned seen = 0;
      for (int a = 0; a < 8; ++a)
      {
        const int o0 = Lx[en[a]] - m[0], o1 = Ly[en[a]] - m[1], o2 = Lz[en[a]] - m[2];
        if (o0 > 1 || o1 > 1 || o2 > 1) return false;
        seen |= 1u << (o0 + 2 * o1 + 4 * o2);
      }
      if (seen != 0xFFu) return false;  // a corner is missing or listed twice
      int& slot = cell2ele[cell(m[0], m[1], m[2])];
      if (slot >= 0) return false;  // two elements in one cell
      slot = e;
    }
    // ---- units over the bounding box of the owned nodes -----------------------------------------------------------
    int lo[3] = {1 << 30, 1 << 30, 1 << 30}, hi[3] = {-1, -1, -1};
    for (int n = 0; n < n_row_nodes; ++n)
    {
      lo[0] = std::min(lo[0], Lx[n]);
      hi[0] = std::max(hi[0], Lx[n]);
      lo[1] = std::min(lo[1], Ly[n]);
      hi[1] = std::max(hi[1], Ly[n]);
      lo[2] = std::min(lo[2], Lz[n]);
      hi[2] = std::max(hi[2], Lz[n]);
    }
    const int ox = hi[0] - lo[0] + 1, oy = hi[1] - lo[1] + 1, oz = hi[2] - lo[2] + 1;
    const int uy = (oy + BY - 1) / BY, uz = (oz + BZ - 1) / BZ;
    if (chunk <= 0)
    {
      // tail of the last wave (1 / (2 waves)) against the extra element layer per chunk (1 / chunk), 148 SMs
      chunk = (int)std::lround(std::sqrt(2.0 * uy * uz * ox / 148.0));
      chunk = std::max(8, chunk);
    }
    chunk = std::min(chunk, ox);
    const int ux = (ox + chunk - 1) / chunk;
    plan.chunk = chunk;

    struct Unit
    {
      int x0, y0, z0, P;
      long long instances;
      bool interior;
    };
    std::vector<Unit> units;
    units.reserve((size_t)ux * uy * uz);
    auto owned_at = [&](int i, int j, int k) -> int {
      if (i < 0 || j < 0 || k < 0 || i >= nx || j >= ny || k >= nz) return -1;
      const int n = lat2node[lat(i, j, k)];
      return (n >= 0 && n < n_row_nodes) ? n : -1;
    };
    // planes are balanced over the chunks of a column (101 planes, 4 chunks: 26 26 25 24 -> 26 25 25 25)
    for (int kz = 0; kz < uz; ++kz)
      for (int ky = 0; ky < uy; ++ky)
        for (int kx = 0; kx < ux; ++kx)
        {
          Unit u;
          u.x0 = lo[0] + (int)((long long)ox * kx / ux);
          u.P = lo[0] + (int)((long long)ox * (kx + 1) / ux) - u.x0;
          u.y0 = lo[1] + ky * BY;
          u.z0 = lo[2] + kz * BZ;
          u.instances = 0;
          u.interior = true;
          if (u.P > 0) units.push_back(u);
        }
    // ---- tables ----------------------------------------------------------------------------------------------
    // first pass: per unit, node table + element table into unit-local vectors, then order the units
    struct UnitData
    {
      std::vector<int> nodes, eles;
      std::vector<unsigned long long> rows;
    };
    std::vector<UnitData> data(units.size());
    for (size_t ui = 0; ui < units.size(); ++ui)
    {
      Unit& u = units[ui];
      UnitData& D = data[ui];
      D.nodes.assign((size_t)u.P * NL, -1);
      bool any = false;
      for (int p = 0; p < u.P; ++p)
        for (int lz = 0; lz < BZ; ++lz)
          for (int ly = 0; ly < BY; ++ly)
          {
            const int n = owned_at(u.x0 + p, u.y0 + ly, u.z0 + lz);
            D.nodes[(size_t)p * NL + ly + BY * lz] = n;
            any = any || n >= 0;
          }
      if (!any)
      {
        u.P = 0;
        continue;
      }
      D.eles.assign((size_t)(u.P + 1) * NE, -1);
      D.rows.assign((size_t)(u.P + 1) * NE, ~0ull);
      for (int t = 0; t <= u.P; ++t)
      {
        const int ex = u.x0 - 1 + t;
        if (ex < 0 || ex >= cx) continue;
        for (int qz = 0; qz < BrickPlanHost::CZ; ++qz)
          for (int qy = 0; qy < CY; ++qy)
          {
            const int ey = u.y0 - 1 + qy, ez = u.z0 - 1 + qz;
            if (ey < 0 || ez < 0 || ey >= cy || ez >= cz) continue;
            const int e = cell2ele[cell(ex, ey, ez)];
            if (e < 0) continue;
            const int* en = conn + (size_t)e * 8;
            unsigned long long rows = 0;
            bool touches = false, ghost = false;
            for (int a = 0; a < 8; ++a)
            {
              const int n = en[a];
              const int p = Lx[n] - u.x0, ly = Ly[n] - u.y0, lz = Lz[n] - u.z0;
              unsigned long long r = 0xFF;
              if (n < n_row_nodes && p >= 0 && p < u.P && ly >= 0 && ly < BY && lz >= 0 && lz < BZ)
              {
                r = (unsigned long long)((p & 1) * NL + ly + BY * lz);
                touches = true;
              }
              ghost = ghost || n >= n_row_nodes;
              rows |= r << (8 * a);
            }
            if (!touches) continue;
            const int k = 12 * ((qy & 1) + 2 * (qz & 1)) + (qy >> 1) + 4 * (qz >> 1);
            D.eles[(size_t)t * NE + k] = e;
            D.rows[(size_t)t * NE + k] = rows;
            u.instances++;
            if (ghost) u.interior = false;
          }
      }
    }
    // interior units first (FC_EVAL_INTERIOR / _BOUNDARY), inside each class the largest first
    std::vector<int> order;
    for (size_t ui = 0; ui < units.size(); ++ui)
      if (units[ui].P > 0) order.push_back((int)ui);
    std::stable_sort(order.begin(), order.end(), [&](int a, int b) {
      if (units[a].interior != units[b].interior) return units[a].interior;
      return units[a].instances > units[b].instances;
    });
    plan.n_units = (int)order.size();
    if (plan.n_units == 0) return false;
    long long noff = 0, eoff = 0;
    for (int ui : order)
    {
      const Unit& u = units[ui];
      plan.unit_planes.push_back(u.P);
      plan.unit_node_off.push_back(noff);
      plan.unit_ele_off.push_back(eoff);
      noff += (long long)u.P * NL;
      eoff += (long long)(u.P + 1) * NE;
      if (u.interior) plan.n_interior_units++;
      plan.n_element_instances += u.instances;
    }
    plan.node_tab.reserve((size_t)noff);
    plan.ele_tab.reserve((size_t)eoff);
    plan.rowidx_tab.reserve((size_t)eoff);
    for (int ui : order)
    {
      UnitData& D = data[ui];
      plan.node_tab.insert(plan.node_tab.end(), D.nodes.begin(), D.nodes.end());
      plan.ele_tab.insert(plan.ele_tab.end(), D.eles.begin(), D.eles.end());
      plan.rowidx_tab.insert(plan.rowidx_tab.end(), D.rows.begin(), D.rows.end());
      std::vector<int>().swap(D.nodes);
      std::vector<int>().swap(D.eles);
      std::vector<unsigned long long>().swap(D.rows);
    }
    return true;
  }
}  // namespace fc_host

// ---- C ABI (include/fourc_b200.h): the plan as plain arrays, for inspection and the CPU test suite -------------
struct fc_brick_plan
{
  fc_host::BrickPlanHost p;
};

extern "C"
{
  int fc_brick_plan_create(int n_col_nodes, const double* X, int n_ele, const int* conn, int n_row_nodes, int max_row_nodes, int chunk,
      fc_brick_plan** out)
  {
    if (!out || !X || !conn || n_col_nodes <= 0 || n_ele < 0 || n_row_nodes < 0) return -1;
    *out = nullptr;
    fc_brick_plan* h = new fc_brick_plan();
    if (!fc_host::build_brick_plan_host(n_col_nodes, X, n_ele, conn, n_row_nodes, max_row_nodes, chunk, h->p))
    {
      delete h;
      return -5;  // FC_ERR_GRAPH: not a lattice of hex cells (the library then scatters element by element)
    }
    *out = h;
    return 0;
  }

  int fc_brick_plan_sizes(const fc_brick_plan* h, int* n_units, int* n_interior_units, int* chunk, int64_t* n_node_entries, int64_t* n_ele_entries,
      int64_t* n_element_instances)
  {
    if (!h) return -1;
    if (n_units) *n_units = h->p.n_units;
    if (n_interior_units) *n_interior_units = h->p.n_interior_units;
    if (chunk) *chunk = h->p.chunk;
    if (n_node_entries) *n_node_entries = (int64_t)h->p.node_tab.size();
    if (n_ele_entries) *n_ele_entries = (int64_t)h->p.ele_tab.size();
    if (n_element_instances) *n_element_instances = h->p.n_element_instances;
    return 0;
  }

  int fc_brick_plan_get(const fc_brick_plan* h, int* unit_planes, int64_t* unit_node_off, int64_t* unit_ele_off, int* node_tab, int* ele_tab,
      uint64_t* rowidx_tab)
  {
    if (!h) return -1;
    const fc_host::BrickPlanHost& p = h->p;
    if (unit_planes) std::copy(p.unit_planes.begin(), p.unit_planes.end(), unit_planes);
    if (unit_node_off) std::copy(p.unit_node_off.begin(), p.unit_node_off.end(), unit_node_off);
    if (unit_ele_off) std::copy(p.unit_ele_off.begin(), p.unit_ele_off.end(), unit_ele_off);
    if (node_tab) std::copy(p.node_tab.begin(), p.node_tab.end(), node_tab);
    if (ele_tab) std::copy(p.ele_tab.begin(), p.ele_tab.end(), ele_tab);
    if (rowidx_tab) std::copy(p.rowidx_tab.begin(), p.rowidx_tab.end(), rowidx_tab);
    return 0;
  }

  void fc_brick_plan_destroy(fc_brick_plan* h) { delete h; }
}
