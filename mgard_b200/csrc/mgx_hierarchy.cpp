// Host-side hierarchy tables.  Mirrors what Hierarchy<D,T,Dev>::init computes on
// the host in the reference (include/mgard-x/Hierarchy/Hierarchy.hpp:22-268 and
// 277-496): level shapes, per-level spacing `dist`, interpolation `ratio`,
// Thomas factors `am/bm` and level volumes, all in the data precision T so the
// kernels see bit-identical constants.  Must be compiled with
// -ffp-contract=off (the build passes it to the host compiler).
#include <cmath>
#include <cstring>

#include "mgx_internal.h"

// Hierarchy.hpp:277-309 -- per dim n -> n/2+1 while n > 2; l_target =
// min_d(#levels) - 1; level_shape[l][d] = shape_level[d][l_target - l].
int mgx_build_levels(int D, const uint64_t *shape, uint64_t max_level,
                     MgxLevels *lv) {
  if (D < 1)
    return MGX_FAILURE;
  if (D > MGX_MAX_D)
    return MGX_TOO_MANY_DIMENSIONS;
  std::vector<std::vector<uint64_t>> sl(D);
  size_t nlevel = SIZE_MAX;
  for (int d = 0; d < D; d++) {
    uint64_t n = shape[d];
    if (n < 3)
      return MGX_FAILURE; // check_shape: no dimension smaller than 3
    while (n > 2) {
      sl[d].push_back(n);
      n = n / 2 + 1;
    }
    sl[d].push_back(2);
    nlevel = std::min(nlevel, sl[d].size());
  }
  uint64_t L = nlevel - 1;
  if (max_level < L)
    L = max_level;
  if (L >= MGX_MAX_L)
    return MGX_FAILURE;
  lv->D = D;
  lv->L = (int)L;
  int pad = MGX_MAX_D - D;
  lv->N = 1;
  for (int d = 0; d < MGX_MAX_D; d++)
    lv->shape[d] = d < pad ? 1 : shape[d - pad];
  uint64_t st = 1;
  for (int d = MGX_MAX_D - 1; d >= 0; d--) {
    lv->stride[d] = st;
    st *= lv->shape[d];
  }
  lv->N = st;
  for (uint64_t l = 0; l <= L; l++)
    for (int d = 0; d < MGX_MAX_D; d++)
      lv->n[l][d] = d < pad ? 1 : sl[d - pad][L - l];
  return MGX_SUCCESS;
}

template <typename T>
void mgx_build_tables(const MgxLevels &lv, MgxHostTables<T> *tb,
                      const void *const *coords, bool normalize) {
  const int L = lv.L;
  const int pad = MGX_MAX_D - lv.D;
  const bool nonuniform = coords != nullptr;
  tb->buf.clear();
  tb->ratios_half = true;
  std::memset(&tb->off, 0, sizeof(tb->off));
  std::vector<std::vector<T>> dist((L + 1) * MGX_MAX_D);
  auto D_ = [&](int l, int d) -> std::vector<T> & {
    return dist[l * MGX_MAX_D + d];
  };
  T vol[MGX_MAX_L][MGX_MAX_D], volr[MGX_MAX_L][MGX_MAX_D];
  for (int d = pad; d < MGX_MAX_D; d++) {
    // create_uniform_coords (0..1 with normalize_coordinates, else 0..n-1) or the
    // caller's node coordinates, then coord_to_dist; Hierarchy.hpp:762-782, 22-61
    uint64_t n = lv.shape[d];
    std::vector<T> coord(n);
    if (nonuniform) {
      const T *c = (const T *)coords[d - pad];
      for (uint64_t i = 0; i < n; i++)
        coord[i] = c[i];
    } else {
      for (uint64_t i = 0; i < n; i++)
        coord[i] = normalize ? (T)i / (n - 1) : (T)i;
    }
    D_(L, d).assign(n, (T)0);
    for (uint64_t i = 0; i + 1 < n; i++)
      D_(L, d)[i] = coord[i + 1] - coord[i];
    // non-uniform grids split the last cell in half when the node count is even
    // (Hierarchy.hpp:42-48); uniform grids keep every level's spacing uniform
    auto split_last = [](std::vector<T> &h) {
      const uint64_t dof = h.size();
      if (dof != 2 && dof % 2 == 0) {
        const T last = h[dof - 2];
        h[dof - 2] = last / (T)2.0;
        h[dof - 1] = last / (T)2.0;
      }
    };
    if (nonuniform)
      split_last(D_(L, d));
    // reduce_dist, Hierarchy.hpp:86-140
    for (int l = L - 1; l >= 0; l--) {
      uint64_t dof = lv.n[l + 1][d], dof2 = lv.n[l][d];
      D_(l, d).assign(dof2, (T)0);
      if (nonuniform) {
        // merge each pair of fine cells, then split the last cell again
        for (uint64_t i = 0; i + 1 < dof2; i++)
          D_(l, d)[i] = D_(l + 1, d)[2 * i] + D_(l + 1, d)[2 * i + 1];
        split_last(D_(l, d));
      } else {
        T total = 0;
        for (uint64_t i = 0; i + 1 < dof; i++)
          total += D_(l + 1, d)[i];
        for (uint64_t i = 0; i + 1 < dof2; i++)
          D_(l, d)[i] = total / (dof2 - 1);
      }
    }
  }
  for (int l = 0; l <= L; l++)
    for (int d = pad; d < MGX_MAX_D; d++) {
      uint64_t dof = lv.n[l][d];
      const std::vector<T> &hd = D_(l, d);
      tb->off.dist[l][d] = (uint32_t)tb->buf.size();
      tb->buf.insert(tb->buf.end(), hd.begin(), hd.end());
      // dist_to_ratio, Hierarchy.hpp:63-84
      tb->off.ratio[l][d] = (uint32_t)tb->buf.size();
      for (uint64_t i = 0; i < dof; i++) {
        T r = 0;
        if (i + 2 < dof) {
          r = hd[i] / (hd[i + 1] + hd[i]);
          if (r != (T)0.5)
            tb->ratios_half = false;
        }
        tb->buf.push_back(r);
      }
      // calc_volume, Hierarchy.hpp:194-268
      T total = 0;
      for (uint64_t i = 0; i + 1 < dof; i++)
        total += hd[i];
      T v = total / (dof - 1);
      vol[l][d] = v;
      volr[l][d] = (T)(1.0 / v);
      // calc_am_bm, Hierarchy.hpp:141-192 (bm stored shifted by one, bm[0]=1)
      std::vector<T> am(dof + 1, (T)0), bm(dof + 1, (T)0), hb(dof + 1, (T)0);
      hb[0] = 2 * hd[0] / 6;
      for (uint64_t i = 1; i + 1 < dof; i++) {
        T a_j = hd[i - 1] / 6;
        T w = a_j / hb[i - 1];
        hb[i] = 2 * (hd[i - 1] + hd[i]) / 6 - w * a_j;
        am[i] = a_j;
      }
      {
        T a_j = hd[dof - 2] / 6;
        T w = a_j / hb[dof - 2];
        hb[dof - 1] = 2 * hd[dof - 2] / 6 - w * a_j;
        am[dof - 1] = a_j;
      }
      bm[0] = 1;
      for (uint64_t i = 0; i < dof; i++)
        bm[i + 1] = hb[i];
      tb->off.am[l][d] = (uint32_t)tb->buf.size();
      tb->buf.insert(tb->buf.end(), am.begin(), am.end());
      tb->off.bm[l][d] = (uint32_t)tb->buf.size();
      tb->buf.insert(tb->buf.end(), bm.begin(), bm.end());
      // am[i] / bm[i], the forward-elimination factor of tridiag_forward2
      // (IPKFunctor.h:13-30); data independent, so rounded once here
      tb->off.abr[l][d] = (uint32_t)tb->buf.size();
      for (uint64_t i = 0; i <= dof; i++)
        tb->buf.push_back(am[i] / bm[i]);
    }
  // lwqzk volume factor, Quantization/LinearQuantization.hpp:74-90: product
  // over d = D-1 .. 0 starting from 1, then sqrt in T
  for (int l = 0; l <= L; l++) {
    T a = 1, b = 1;
    for (int d = MGX_MAX_D - 1; d >= pad; d--) {
      a *= vol[l][d];
      b *= volr[l][d];
    }
    tb->vol[l] = std::sqrt(a);
    tb->volr[l] = std::sqrt(b);
  }
}

template void mgx_build_tables<float>(const MgxLevels &, MgxHostTables<float> *,
                                      const void *const *, bool);
template void mgx_build_tables<double>(const MgxLevels &, MgxHostTables<double> *,
                                       const void *const *, bool);

// ---- host-only entry points of the C ABI ---------------------------------------
#include <cmath>
#include <limits>

template <typename T>
static T local_abs_tol_t(int rel, T norm, T tol, T s, uint64_t num_subdomain) {
  if (rel) {
    if (s == std::numeric_limits<T>::infinity())
      return tol * norm;
    return std::sqrt((tol * norm) * (tol * norm) / num_subdomain);
  }
  if (s == std::numeric_limits<T>::infinity())
    return tol;
  return std::sqrt((tol * tol) / num_subdomain);
}

extern "C" {

int mgx_hierarchy_levels(int D, const uint64_t *shape, uint64_t max_level,
                         int *l_target, uint64_t *level_shapes) {
  if (D > MGX_MAX_D)
    return MGX_TOO_MANY_DIMENSIONS;
  if (D < 1 || !shape)
    return MGX_FAILURE;
  MgxLevels lv;
  int rc = mgx_build_levels(D, shape, max_level, &lv);
  if (rc)
    return rc;
  const int pad = MGX_MAX_D - D;
  if (l_target)
    *l_target = lv.L;
  if (level_shapes)
    for (int l = 0; l <= lv.L; l++)
      for (int d = 0; d < D; d++)
        level_shapes[l * D + d] = lv.n[l][d + pad];
  return MGX_SUCCESS;
}

int mgx_max_dim_decomposition(int D, const uint64_t *shape,
                              uint64_t min_subdomains, int *dim, uint64_t *size,
                              uint64_t *num) {
  if (D > MGX_MAX_D)
    return MGX_TOO_MANY_DIMENSIONS;
  if (D < 1 || !shape || !dim || !size || !num)
    return MGX_FAILURE;
  uint64_t best = 0;
  int bd = 0;
  for (int d = 0; d < D; d++)
    if (shape[d] > best) {
      best = shape[d];
      bd = d;
    }
  uint64_t sz = best, n = 1;
  while (n < min_subdomains && sz > 1) {
    sz = (sz - 1) / 2 + 1; // divide by 2 and round up
    n = (best - 1) / sz + 1;
  }
  *dim = bd;
  *size = sz;
  *num = n;
  return MGX_SUCCESS;
}

int mgx_subdomain_shape(int D, const uint64_t *shape, int dim, uint64_t size,
                        uint64_t id, uint64_t *sub_shape, uint64_t *offset) {
  if (D < 1 || D > MGX_MAX_D || !shape || dim < 0 || dim >= D || size == 0)
    return MGX_FAILURE;
  const uint64_t num = (shape[dim] - 1) / size + 1;
  if (id >= num)
    return MGX_FAILURE;
  for (int d = 0; d < D; d++)
    sub_shape[d] = shape[d];
  // full slabs first, then the leftover (DomainDecomposer.hpp:125-133)
  if (id < shape[dim] / size)
    sub_shape[dim] = size;
  else
    sub_shape[dim] = shape[dim] % size;
  if (offset)
    *offset = id * size;
  return MGX_SUCCESS;
}

double mgx_local_abs_tol(int dtype, int rel, double norm, double tol, double s,
                         uint64_t num_subdomains) {
  if (dtype == MGX_F32)
    return (double)local_abs_tol_t<float>(rel, (float)norm, (float)tol, (float)s,
                                          num_subdomains);
  return local_abs_tol_t<double>(rel, norm, tol, s, num_subdomains);
}

} // extern "C"
