// plan.cpp -- the one-time integer pass: sparsity pattern + element -> slot map.
//
// Replaces what PETSc derives implicitly from the reference's MatSetValues(ADD_VALUES)
// calls (src/main.F90:157-171): row r of the assembled AIJ holds, in ascending order, every
// global dof coupled to r through some element -- explicit zeros included, so cohesive
// elements (Ke = 0, src/m_global.F90:250-252) still contribute to the pattern.  Because a
// node's pdim dofs are consecutive columns and local node ids ascend with global ids
// (src/main.F90:57-63), the pattern is a node-block CSR; scalar slots are closed-form.
#include <algorithm>
#include <chrono>
#include <thread>

#include "femx_internal.h"

namespace femx {

namespace {
template <class F>
void parallel_for(int64_t n, F f) {
  int nt = (int)std::thread::hardware_concurrency();
  if (nt < 1) nt = 1;
  if (nt > 32) nt = 32;
  if (n < 4096) nt = 1;
  if (nt == 1) { f(0, n); return; }
  std::vector<std::thread> th;
  for (int t = 0; t < nt; ++t) th.emplace_back(f, n * t / nt, n * (t + 1) / nt);
  for (auto &x : th) x.join();
}
}  // namespace

int build_plan(int pdim, int nels, const int *etype, const int *eptr, const int *eind,
               const int *enlmat, int nlnds, int nnds, Plan *P, std::string *err) {
  auto t0 = std::chrono::steady_clock::now();
  P->pdim = pdim; P->nels = nels; P->nlnds = nlnds; P->nnds = nnds;
  const int64_t nconn = eptr[nels];
  P->conn0.resize(nconn);
  for (int64_t i = 0; i < nconn; ++i) {
    const int v = eind[i] - 1;
    if (v < 0 || v >= nlnds) { *err = "element node id out of range (expects 1-based local ids)"; return FEMX_ERR_ARG; }
    P->conn0[i] = v;
  }
  P->eb_ptr.assign(nels + 1, 0); P->ke_off.assign(nels + 1, 0); P->fe_off.assign(nels + 1, 0);
  for (int e = 0; e < nels; ++e) {
    const int t = etype[e];
    if (t < FEMX_TRI || t > FEMX_COH) { *err = "unknown element type code"; return FEMX_ERR_ARG; }
    const int nen = type_nen(t);
    if (eptr[e + 1] - eptr[e] != nen) { *err = "eptr does not match the element type's node count"; return FEMX_ERR_ARG; }
    if (type_dim(t) != pdim) { *err = "element type dimension differs from pdim"; return FEMX_ERR_ARG; }
    P->eb_ptr[e + 1] = P->eb_ptr[e] + nen * nen;
    P->ke_off[e + 1] = P->ke_off[e] + (int64_t)nen * nen * pdim * pdim;
    P->fe_off[e + 1] = P->fe_off[e] + nen * pdim;
  }
  // node -> (element, a), ascending element (counting sort over the element loop)
  P->n2e_ptr.assign(nlnds + 1, 0);
  for (int64_t i = 0; i < nconn; ++i) P->n2e_ptr[P->conn0[i] + 1]++;
  for (int r = 0; r < nlnds; ++r) P->n2e_ptr[r + 1] += P->n2e_ptr[r];
  P->n2e_elem.resize(nconn); P->n2e_a.resize(nconn);
  {
    std::vector<int64_t> fill(P->n2e_ptr.begin(), P->n2e_ptr.end() - 1);
    for (int e = 0; e < nels; ++e)
      for (int a = 0; a < eptr[e + 1] - eptr[e]; ++a) {
        const int r = P->conn0[eptr[e] + a];
        P->n2e_elem[fill[r]] = e; P->n2e_a[fill[r]] = (uint8_t)a; ++fill[r];
      }
  }
  // block rows: union of the nodes of the incident elements, ascending
  P->bptr.assign(nlnds + 1, 0);
  std::vector<int32_t> cnt(nlnds, 0);
  auto row_nodes = [&](int r, std::vector<int32_t> &buf) {
    buf.clear();
    for (int64_t p = P->n2e_ptr[r]; p < P->n2e_ptr[r + 1]; ++p) {
      const int e = P->n2e_elem[p];
      for (int k = eptr[e]; k < eptr[e + 1]; ++k) buf.push_back(P->conn0[k]);
    }
    std::sort(buf.begin(), buf.end());
    buf.erase(std::unique(buf.begin(), buf.end()), buf.end());
  };
  parallel_for(nlnds, [&](int64_t lo, int64_t hi) {
    std::vector<int32_t> buf;
    for (int64_t r = lo; r < hi; ++r) { row_nodes((int)r, buf); cnt[r] = (int32_t)buf.size(); }
  });
  for (int r = 0; r < nlnds; ++r) {
    if (cnt[r] > 65535) { *err = "a node couples to more than 65535 nodes"; return FEMX_ERR_UNSUPPORTED; }
    P->bptr[r + 1] = P->bptr[r] + cnt[r];
  }
  P->nblk = P->bptr[nlnds];
  P->bcol.resize(P->nblk); P->bdiag.assign(nlnds, 0);
  parallel_for(nlnds, [&](int64_t lo, int64_t hi) {
    std::vector<int32_t> buf;
    for (int64_t r = lo; r < hi; ++r) {
      row_nodes((int)r, buf);
      std::copy(buf.begin(), buf.end(), P->bcol.begin() + P->bptr[r]);
      P->bdiag[r] = (int32_t)(std::lower_bound(buf.begin(), buf.end(), (int32_t)r) - buf.begin());
    }
  });
  // element -> block position
  P->bpos.resize(P->eb_ptr[nels]);
  parallel_for(nels, [&](int64_t lo, int64_t hi) {
    for (int64_t e = lo; e < hi; ++e) {
      const int nen = eptr[e + 1] - eptr[e];
      for (int a = 0; a < nen; ++a) {
        const int r = P->conn0[eptr[e] + a];
        const int32_t *b0 = &P->bcol[P->bptr[r]], *b1 = &P->bcol[P->bptr[r + 1]];
        for (int b = 0; b < nen; ++b)
          P->bpos[P->eb_ptr[e] + a * nen + b] = (uint16_t)(std::lower_bound(b0, b1, P->conn0[eptr[e] + b]) - b0);
      }
    }
  });
  // row gather lists (fused rows kernel): invert bpos per node
  {
    P->rbe_ptr.assign(nlnds + 1, 0);
    P->max_inc = 0;
    for (int r = 0; r < nlnds; ++r) {
      int64_t n = 0;
      for (int64_t p = P->n2e_ptr[r]; p < P->n2e_ptr[r + 1]; ++p) { const int e = P->n2e_elem[p]; n += eptr[e + 1] - eptr[e]; }
      P->rbe_ptr[r + 1] = P->rbe_ptr[r] + n;
      P->max_inc = std::max(P->max_inc, (int)(P->n2e_ptr[r + 1] - P->n2e_ptr[r]));
    }
    int maxnen = 3;
    for (int e = 0; e < nels; ++e) maxnen = std::max(maxnen, eptr[e + 1] - eptr[e]);
    const int d2 = pdim * pdim;
    P->rb_foff = maxnen * d2;
    P->rb_slot = P->rb_foff + pdim;
    if (P->rb_slot % 2 == 0) ++P->rb_slot;   // odd pitch: lane-strided 8-byte stores hit distinct banks
    const int slot = P->rb_slot;
    P->rb_off.assign((size_t)P->nblk + nlnds, 0);
    P->rb_ent.assign((size_t)P->rbe_ptr[nlnds], 0);
    P->simplex = nels > 0;
    for (int e = 0; e < nels; ++e) P->simplex = P->simplex && (etype[e] == FEMX_TRI || etype[e] == FEMX_TET);
    P->sx_ent.assign(P->simplex ? (size_t)P->rbe_ptr[nlnds] : 0, 0);
    P->sx_inc.assign(P->simplex ? 4 * (size_t)P->n2e_ptr[nlnds] : 0, 0);
    P->sx_mat.assign(P->simplex ? (size_t)P->n2e_ptr[nlnds] : 0, 0);
    if (P->max_inc <= 32)
      parallel_for(nlnds, [&](int64_t lo, int64_t hi) {
        std::vector<int32_t> cntk;
        for (int64_t r = lo; r < hi; ++r) {
          const int nb = (int)(P->bptr[r + 1] - P->bptr[r]);
          cntk.assign(nb + 1, 0);
          for (int64_t p = P->n2e_ptr[r]; p < P->n2e_ptr[r + 1]; ++p) {
            const int e = P->n2e_elem[p], a = P->n2e_a[p], nen = eptr[e + 1] - eptr[e];
            for (int b = 0; b < nen; ++b) ++cntk[P->bpos[P->eb_ptr[e] + a * nen + b] + 1];
          }
          for (int k = 0; k < nb; ++k) cntk[k + 1] += cntk[k];
          uint16_t *off = &P->rb_off[P->bptr[r] + r];
          for (int k = 0; k <= nb; ++k) off[k] = (uint16_t)cntk[k];
          uint16_t *ent = &P->rb_ent[P->rbe_ptr[r]];
          int idx = 0;
          for (int64_t p = P->n2e_ptr[r]; p < P->n2e_ptr[r + 1]; ++p, ++idx) {   // ascending element order
            const int e = P->n2e_elem[p], a = P->n2e_a[p], nen = eptr[e + 1] - eptr[e];
            if (P->simplex) {   // flattened incidence record: no element -> connectivity -> node chain on the device
              for (int b = 0; b < nen; ++b) P->sx_inc[4 * p + b] = P->conn0[eptr[e] + b];
            }
            for (int b = 0; b < nen; ++b) {
              const int c = cntk[P->bpos[P->eb_ptr[e] + a * nen + b]]++;
              ent[c] = (uint16_t)(idx * slot + b * d2);
              if (P->simplex) P->sx_ent[P->rbe_ptr[r] + c] = (uint16_t)((idx << 4) | (a << 2) | b);
            }
          }
        }
      });
  }
  // cohesive overlay (elements with nlMat != 0, src/m_global.F90:130-150, :188-206)
  P->coh_elems.clear();
  for (int e = 0; e < nels; ++e)
    if (enlmat[e] != 0) {
      if (etype[e] != FEMX_COH) { *err = "nlMat != 0 on a non-cohesive element"; return FEMX_ERR_UNSUPPORTED; }
      P->coh_elems.push_back(e);
    }
  {
    struct C { int64_t blk; int32_t row, src; };
    std::vector<C> cs;
    struct N { int32_t node, src; };
    std::vector<N> ns;
    for (size_t ci = 0; ci < P->coh_elems.size(); ++ci) {
      const int e = P->coh_elems[ci];
      for (int a = 0; a < 4; ++a) {
        const int r = P->conn0[eptr[e] + a];
        ns.push_back({r, (int32_t)((ci << 4) | a)});
        for (int b = 0; b < 4; ++b)
          cs.push_back({P->bptr[r] + P->bpos[P->eb_ptr[e] + a * 4 + b], r, (int32_t)((ci << 8) | (a << 4) | b)});
      }
    }
    std::stable_sort(cs.begin(), cs.end(), [](const C &x, const C &y) { return x.blk < y.blk; });
    std::stable_sort(ns.begin(), ns.end(), [](const N &x, const N &y) { return x.node < y.node; });
    P->cblk_id.clear(); P->cblk_row.clear(); P->cblk_ptr.assign(1, 0); P->cblk_src.clear();
    for (size_t i = 0; i < cs.size(); ++i) {
      if (i == 0 || cs[i].blk != cs[i - 1].blk) {
        if (i) P->cblk_ptr.push_back((int32_t)i);
        P->cblk_id.push_back(cs[i].blk); P->cblk_row.push_back(cs[i].row);
      }
      P->cblk_src.push_back(cs[i].src);
    }
    if (!cs.empty()) P->cblk_ptr.push_back((int32_t)cs.size());
    P->cnode_id.clear(); P->cnode_ptr.assign(1, 0); P->cnode_src.clear();
    for (size_t i = 0; i < ns.size(); ++i) {
      if (i == 0 || ns[i].node != ns[i - 1].node) {
        if (i) P->cnode_ptr.push_back((int32_t)i);
        P->cnode_id.push_back(ns[i].node);
      }
      P->cnode_src.push_back(ns[i].src);
    }
    if (!ns.empty()) P->cnode_ptr.push_back((int32_t)ns.size());
  }
  P->build_seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
  return FEMX_OK;
}

}  // namespace femx
