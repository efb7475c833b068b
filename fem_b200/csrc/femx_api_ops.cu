// femx_api_ops.cu -- element-level and mesh-level entry points of include/femx.h.
#include "api_util.h"

namespace {

// persistent device buffer, grown on demand
int ensure(femx_ctx *c, double **p, int64_t n) {
  auto it = c->buf_n.find(p);
  if (*p && it != c->buf_n.end() && it->second >= n) return FEMX_OK;
  if (*p) { dev_release(c, *p); *p = nullptr; }
  c->buf_n[p] = n;
  return dev_alloc(c, p, n);
}

// device pointer for an input vector: the caller's (DEVICE) or a persistent staging copy
int stage_in(femx_ctx *c, const double *src, int64_t n, int where, double **slot, const double **out) {
  if (where == FEMX_DEVICE) { *out = src; return FEMX_OK; }
  int rc = ensure(c, slot, n);
  if (rc) return rc;
  CK(cudaMemcpyAsync(*slot, src, sizeof(double) * (size_t)n, cudaMemcpyHostToDevice, c->stream));
  *out = *slot;
  return FEMX_OK;
}

int stage_out(femx_ctx *c, double *dst, int64_t n, int where, double **slot, double **out) {
  if (where == FEMX_DEVICE) { *out = dst; return FEMX_OK; }
  int rc = ensure(c, slot, n);
  if (rc) return rc;
  *out = *slot;
  return FEMX_OK;
}

int copy_back(femx_ctx *c, double *dst, const double *dev, int64_t n, int where) {
  if (where == FEMX_DEVICE || dst == dev) return FEMX_OK;
  CK(cudaMemcpyAsync(dst, dev, sizeof(double) * (size_t)n, cudaMemcpyDeviceToHost, c->stream));
  return FEMX_OK;
}

int assemble(femx_ctx *c, const double *u_local, double *vals, double *fint, int where) {
  ARG(c->has_plan, "femx_build_plan first");
  ARG(vals, "vals is NULL");
  ARG((u_local == nullptr) == (fint == nullptr), "u_local and fint go together");
  CK(cudaSetDevice(c->device));
  const int64_t nloc = (int64_t)c->pdim * c->nlnds, nglob = (int64_t)c->pdim * c->nnds;
  int rc;
  const double *d_u = nullptr;
  double *d_vals = nullptr, *d_fint = nullptr;
  if (u_local && (rc = stage_in(c, u_local, nloc, where, &c->d_u, &d_u))) return rc;
  if ((rc = stage_out(c, vals, c->nnz, where, &c->d_vals, &d_vals))) return rc;
  if (fint && (rc = stage_out(c, fint, nglob, where, &c->d_vec[0], &d_fint))) return rc;
  // auto (0): the warp-specialised DMMA tile kernel (20) for a logical hex8 grid (the symmetric pair-block kernel 12 when the assembly
  // kernel itself is to store the interface rows at their owners), else the simplex / fused rows kernels, else two-pass
  const int av0 = c->opts.assembly_variant;
  // the hex8 tile kernels address values with 32-bit slots: beyond 2^32 values (34 GB, ~17.5 M hex8 on one GPU) auto uses the general kernels
  const bool slots32 = c->nnz <= 0xffffffffll;
  const bool fused = c->hex.ok && c->mat_ok_hex && slots32 && av0 != 1 && av0 != 9 && av0 != 11;
  if (!fused && (av0 == 12 || av0 == 13 || av0 == 20 || av0 == 21)) {
    c->err = "hex8 tile kernel: " + (!c->hex.ok ? c->hex.why : !c->mat_ok_hex ? c->mat_why : std::string("more than 2^32 matrix values"));
    return FEMX_ERR_UNSUPPORTED;
  }
  const bool simplex = !fused && av0 != 1 && av0 != 9 && c->mat_ok_simplex && rows_simplex_ok(c);   // pure tri3 / tet4 mesh
  const bool rows = !fused && !simplex && av0 != 1 && rows_fused_ok(c);
  if (av0 == 9 && !rows) { c->err = "fused rows kernel: more than 32 elements around a node"; return FEMX_ERR_UNSUPPORTED; }
  if (av0 == 11 && !simplex) { c->err = "simplex rows kernel: needs a pure tri3 / tet4 mesh with at most 32 elements around a node"; return FEMX_ERR_UNSUPPORTED; }
  const bool peer = c->peer.connected;
  // peer exchange: reserved[3] = 1 lets the assembly kernel store interface rows at their owner itself (variants 0, 5, 12, 9);
  // default is the push kernel behind the unchanged assembly kernel
  const bool peer_fused = peer && c->opts.reserved[3] == 1 && (fused ? (av0 == 0 || av0 == 12) : (rows || simplex));
  c->peer.in_kernel = peer_fused;
  if (!fused && !rows && !simplex) {
    if ((rc = ensure(c, &c->d_ke, c->plan.ke_off[c->nels]))) return rc;
    if (fint && (rc = ensure(c, &c->d_fe, c->plan.fe_off[c->nels]))) return rc;
  }
  if ((rc = begin_timing(c))) return rc;
  if (fint && c->nlnds != c->nnds) CK(cudaMemsetAsync(d_fint, 0, sizeof(double) * (size_t)nglob, c->stream));
  bool split = false;
  if (fused) {
    const int av = av0 == 0 ? (peer_fused ? 12 : 20) : av0;
#ifdef FEMX_DIAG
    if (av == 13) c->launches += launch_hex_sym7(c, d_u, d_vals, d_fint);
    else if (av == 21) c->launches += launch_hex_mma_prof(c, d_u, d_vals, d_fint);
    else
#endif
    // peer exchange behind variant 20: interface CTAs first on a high-priority stream, the exchange behind them, interior CTAs
    // meanwhile on the ctx stream (FEMX_PEER_OVERLAP=0 keeps one launch and the exchange behind it)
    static const bool overlap_on = !(getenv("FEMX_PEER_OVERLAP") && atoi(getenv("FEMX_PEER_OVERLAP")) == 0);
    // (the interface rows are found through the force lists -- every node another rank holds sends or receives a force entry; a host
    // that set a matrix merge list without a force merge list gets the serial order)
    if (av == 20 && peer && !peer_fused && overlap_on && (c->peer.n_fmerge > 0 || c->peer.n_merge == 0)) {
      if ((rc = peer_hex_split(c))) return rc;
      split = c->peer.n_cta_if > 0 && c->peer.n_cta_if < c->peer.n_cta;
    }
    if (split) {
      auto &P = c->peer;
      CK(cudaEventRecord(P.ev_fork, c->stream));
      CK(cudaStreamWaitEvent(P.hi, P.ev_fork, 0));
      c->launches += launch_hex_mma_part(c, d_u, d_vals, d_fint, P.d_cta_list, P.n_cta_if, P.hi);
      c->launches += launch_hex_mma_part(c, d_u, d_vals, d_fint, P.d_cta_list + P.n_cta_if, P.n_cta - P.n_cta_if, c->stream);
    } else
    c->launches += av == 20 ? launch_hex_mma(c, d_u, d_vals, d_fint) : launch_hex_sym6(c, d_u, d_vals, d_fint);
  } else if (simplex) {
    c->launches += launch_rows_simplex(c, d_u, d_vals, d_fint);
  } else if (rows) {
    c->launches += launch_rows_fused(c, d_u, d_vals, d_fint);
  } else {
    c->launches += launch_elem_rows(c, d_u, c->d_ke, fint ? c->d_fe : nullptr);
    c->launches += launch_gather_rows(c, c->d_ke, fint ? c->d_fe : nullptr, d_vals, d_fint);
  }
  c->mid_valid = false;
  if (peer) {
    if (!c->ev_mid) CK(cudaEventCreate(&c->ev_mid));
    CK(cudaEventRecord(c->ev_mid, c->stream));
    c->mid_valid = true;
    if (split) {   // the exchange follows the interface CTAs on their stream; the ctx stream joins when both are done
      auto &P = c->peer;
      cudaStream_t main = c->stream;
      c->stream = P.hi;
      const int nl = launch_peer_finish(c, peer_fused, d_vals, d_fint);
      c->stream = main;
      c->launches += nl;
      CK(cudaEventRecord(P.ev_join, P.hi));
      CK(cudaStreamWaitEvent(c->stream, P.ev_join, 0));
    } else
    c->launches += launch_peer_finish(c, peer_fused, d_vals, d_fint);   // push, step barrier (the peers' stores into this rank's arena are complete), receiver add
  }
  if ((rc = end_timing(c))) return rc;
  if ((rc = copy_back(c, vals, d_vals, c->nnz, where))) return rc;
  if (fint && (rc = copy_back(c, fint, d_fint, nglob, where))) return rc;
  if ((rc = finish_call(c, true))) return rc;
  if (peer && !c->async && (rc = femx_peer_last_error(c))) return rc;
  return FEMX_OK;
}

}  // namespace

extern "C" {

// --------------------------------------------------------------------------- element level
int femx_element_ke(femx_ctx *c, int type, int n, const double *ecoords, const double *E_nu, double *ke) {
  if (!c) return FEMX_ERR_ARG;
  ARG(type >= FEMX_TRI && type <= FEMX_HEX, "femx_element_ke: type must be tri/qua/tet/hex");
  ARG(n >= 0 && ecoords && E_nu && ke, "bad arguments");
  if (n == 0) return FEMX_OK;
  CK(cudaSetDevice(c->device));
  const int nen = type_nen(type), dim = type_dim(type), nd = nen * dim;
  View x(c, ecoords, sizeof(double) * (size_t)n * nen * dim, FEMX_HOST, true, false);
  View m(c, E_nu, sizeof(double) * (size_t)n * 2, FEMX_HOST, true, false);
  View k(c, ke, sizeof(double) * (size_t)n * nd * nd, FEMX_HOST, false, true);
  if (x.rc || m.rc || k.rc) return x.rc | m.rc | k.rc;
  c->launches += launch_element_ke(c, type, n, x.as<double>(), m.as<double>(), k.as<double>());
  int rc = k.finish();
  if (rc) return rc;
  return finish_call(c, false);
}

int femx_element_stress(femx_ctx *c, int type, int n, const double *ecoords, const double *edisp,
                        const double *E_nu, double *estress, double *enodal) {
  if (!c) return FEMX_ERR_ARG;
  ARG(type >= FEMX_TRI && type <= FEMX_HEX, "femx_element_stress: type must be tri/qua/tet/hex");
  ARG(n >= 0 && ecoords && edisp && E_nu && estress, "bad arguments");
  if (n == 0) return FEMX_OK;
  CK(cudaSetDevice(c->device));
  const int nen = type_nen(type), dim = type_dim(type), nip = type_nip(type), cp = dim == 2 ? 3 : 6;
  View x(c, ecoords, sizeof(double) * (size_t)n * nen * dim, FEMX_HOST, true, false);
  View u(c, edisp, sizeof(double) * (size_t)n * nen * dim, FEMX_HOST, true, false);
  View m(c, E_nu, sizeof(double) * (size_t)n * 2, FEMX_HOST, true, false);
  View s(c, estress, sizeof(double) * (size_t)n * nip * cp, FEMX_HOST, false, true);
  View sn(c, enodal, sizeof(double) * (size_t)n * nen * cp, FEMX_HOST, false, true);
  if (x.rc || u.rc || m.rc || s.rc || sn.rc) return FEMX_ERR_ALLOC;
  c->launches += launch_element_stress(c, type, n, x.as<double>(), u.as<double>(), m.as<double>(), s.as<double>(),
                                       sn.as<double>());
  int rc = s.finish();
  if (rc) return rc;
  if ((rc = sn.finish())) return rc;
  return finish_call(c, false);
}

int femx_element_cohesive(femx_ctx *c, int n, const double *ecoords, const int *nodes, const double *u_local,
                          int nu_local, const double *props, double dt, double *force, double *stiff) {
  if (!c) return FEMX_ERR_ARG;
  ARG(n >= 0 && ecoords && nodes && u_local && props && force && stiff, "bad arguments");
  ARG(nu_local >= 8, "u_local must hold at least 4 nodes");
  if (n == 0) return FEMX_OK;
  CK(cudaSetDevice(c->device));
  View x(c, ecoords, sizeof(double) * (size_t)n * 8, FEMX_HOST, true, false);
  View nd(c, nodes, sizeof(int) * (size_t)n * 4, FEMX_HOST, true, false);
  View u(c, u_local, sizeof(double) * (size_t)nu_local, FEMX_HOST, true, false);
  View p(c, props, sizeof(double) * (size_t)n * 6, FEMX_HOST, true, false);
  View f(c, force, sizeof(double) * (size_t)n * 8, FEMX_HOST, false, true);
  View s(c, stiff, sizeof(double) * (size_t)n * 64, FEMX_HOST, false, true);
  if (x.rc || nd.rc || u.rc || p.rc || f.rc || s.rc) return FEMX_ERR_ALLOC;
  c->launches += launch_cohesive_unit(c, n, x.as<double>(), nd.as<int32_t>(), u.as<double>(), p.as<double>(), dt,
                                      f.as<double>(), s.as<double>());
  int rc = f.finish();
  if (rc) return rc;
  if ((rc = s.finish())) return rc;
  return finish_call(c, false);
}

// --------------------------------------------------------------------------- mesh level
int femx_assemble_K(femx_ctx *c, double *vals, int where) {
  if (!c) return FEMX_ERR_ARG;
  return assemble(c, nullptr, vals, nullptr, where);
}

int femx_assemble_K_fint(femx_ctx *c, const double *u_local, double *vals, double *fint, int where) {
  if (!c) return FEMX_ERR_ARG;
  ARG(u_local && fint, "u_local / fint is NULL");
  return assemble(c, u_local, vals, fint, where);
}

int femx_apply_kbc(femx_ctx *c, double *vals, int where) {
  if (!c) return FEMX_ERR_ARG;
  ARG(c->has_plan && vals, "femx_build_plan first / vals NULL");
  CK(cudaSetDevice(c->device));
  if (where == FEMX_DEVICE) {
    c->launches += launch_apply_kbc(c, vals);
    return finish_call(c, false);
  }
  // host values: the BC list is O(#BC); patch through a staged copy of the touched slots
  View v(c, vals, sizeof(double) * (size_t)c->nnz, FEMX_HOST, true, true);
  if (v.rc) return v.rc;
  c->launches += launch_apply_kbc(c, v.as<double>());
  int rc = v.finish();
  if (rc) return rc;
  return finish_call(c, false);
}

static int jacobian_impl(femx_ctx *c, const double *kvals, const double *u_total_local, double dt, double *jvals, int where, bool update) {
  ARG(c->has_plan && kvals && u_total_local && jvals, "femx_build_plan first / NULL argument");
  ARG(!update || (where == FEMX_DEVICE && jvals != kvals), "femx_jacobian_update works on a resident Jacobian (FEMX_DEVICE, jvals != kvals)");
  CK(cudaSetDevice(c->device));
  int rc;
  const double *d_u = nullptr, *d_k = nullptr;
  double *d_j = nullptr;
  if ((rc = stage_in(c, u_total_local, (int64_t)c->pdim * c->nlnds, where, &c->d_u, &d_u))) return rc;
  if ((rc = stage_in(c, kvals, c->nnz, where, &c->d_kstage, &d_k))) return rc;
  if ((rc = stage_out(c, jvals, c->nnz, where, &c->d_vals, &d_j))) return rc;
  if ((rc = begin_timing(c))) return rc;
  // MatCopy(Mat_K, Jacobian) src/m_global.F90:123; the update form rewrites only what a cohesive element or a BC touches
  if (d_j != d_k && !update) CK(cudaMemcpyAsync(d_j, d_k, sizeof(double) * (size_t)c->nnz, cudaMemcpyDeviceToDevice, c->stream));
  c->launches += launch_cohesive_mesh(c, d_u, dt, nullptr, c->d_coh_ke);
  c->launches += launch_jacobian_overlay(c, d_k, d_j);
  c->launches += launch_apply_kbc(c, d_j);  // ApplyKBC(Jacobian) :154
  if ((rc = end_timing(c))) return rc;
  if ((rc = copy_back(c, jvals, d_j, c->nnz, where))) return rc;
  return finish_call(c, true);
}

int femx_jacobian(femx_ctx *c, const double *kvals, const double *u_total_local, double dt, double *jvals, int where) {
  if (!c) return FEMX_ERR_ARG;
  return jacobian_impl(c, kvals, u_total_local, dt, jvals, where, false);
}

int femx_jacobian_update(femx_ctx *c, const double *kvals, const double *u_total_local, double dt, double *jvals, int where) {
  if (!c) return FEMX_ERR_ARG;
  return jacobian_impl(c, kvals, u_total_local, dt, jvals, where, true);
}

static int residual_impl(femx_ctx *c, const double *jvals, const double *u_total_local, const double *du,
                         const double *f_ext, double dt, double t_total, double *r, int where) {
  ARG(c->has_plan && u_total_local && f_ext && r, "femx_build_plan first / NULL argument");
  ARG(t_total != 0.0, "t_total is zero");
  CK(cudaSetDevice(c->device));
  const int64_t nglob = (int64_t)c->pdim * c->nnds;
  int rc;
  const double *d_u = nullptr, *d_f = nullptr, *d_du = nullptr, *d_j = nullptr;
  double *d_r = nullptr, *d_br = nullptr;
  if ((rc = stage_in(c, u_total_local, (int64_t)c->pdim * c->nlnds, where, &c->d_u, &d_u))) return rc;
  if ((rc = stage_in(c, f_ext, nglob, where, &c->d_vec[1], &d_f))) return rc;
  if ((rc = stage_out(c, r, nglob, where, &c->d_vec[0], &d_r))) return rc;
  if (jvals) {
    if ((rc = stage_in(c, du, nglob, where, &c->d_vec[2], &d_du))) return rc;
    if ((rc = stage_in(c, jvals, c->nnz, where, &c->d_vals, &d_j))) return rc;
    // bracket goes to a scratch vector, r gets J*du - bracket
    if ((rc = ensure(c, &c->d_vec[3], nglob))) return rc;
    d_br = c->d_vec[3];
  } else {
    d_br = d_r;
  }
  if ((rc = begin_timing(c))) return rc;
  CK(cudaMemcpyAsync(d_br, d_f, sizeof(double) * (size_t)nglob, cudaMemcpyDeviceToDevice, c->stream));  // VecCopy(Vec_F, R)
  c->launches += launch_cohesive_mesh(c, d_u, dt, c->d_coh_fe, nullptr);
  c->launches += launch_cohesive_force_gather(c, d_br);
  c->launches += launch_bc_force(c, dt, t_total, d_br);  // EnforceBCForce :439-455
  if (jvals) {
    if (c->nlnds != c->nnds) CK(cudaMemsetAsync(d_r, 0, sizeof(double) * (size_t)nglob, c->stream));
    c->launches += launch_spmv_sub(c, d_j, d_du, d_br, d_r);  // VecScale(-1); MatMultAdd :225-226
  }
  if ((rc = end_timing(c))) return rc;
  if ((rc = copy_back(c, r, d_r, nglob, where))) return rc;
  return finish_call(c, true);
}

int femx_residual(femx_ctx *c, const double *jvals, const double *u_total_local, const double *du,
                  const double *f_ext, double dt, double t_total, double *r, int where) {
  if (!c) return FEMX_ERR_ARG;
  ARG(jvals && du, "jvals / du is NULL");
  return residual_impl(c, jvals, u_total_local, du, f_ext, dt, t_total, r, where);
}

int femx_residual_forces(femx_ctx *c, const double *u_total_local, const double *f_ext, double dt, double t_total,
                         double *out, int where) {
  if (!c) return FEMX_ERR_ARG;
  return residual_impl(c, nullptr, u_total_local, nullptr, f_ext, dt, t_total, out, where);
}

// ---- loads: ReadDistForces / ReadDistTractions arrays (src/m_io.F90:330-372, :414-475) -> FormRHS
int femx_set_loads(femx_ctx *c, int nfrcs, const int *fnode, const double *fval, int ntrcs, const int *telsd, const double *tval) {
  if (!c) return FEMX_ERR_ARG;
  ARG(c->has_plan, "femx_build_plan first");
  ARG(nfrcs >= 0 && ntrcs >= 0 && (nfrcs == 0 || (fnode && fval)) && (ntrcs == 0 || (telsd && tval)), "bad load arrays");
  CK(cudaSetDevice(c->device));
  const int pd = c->pdim, w = pd + 2;
  static const int tri[3][2] = {{0, 1}, {1, 2}, {2, 0}}, qua[4][2] = {{0, 1}, {1, 2}, {2, 3}, {3, 0}};
  static const int tet[4][3] = {{0, 1, 3}, {1, 2, 3}, {0, 2, 3}, {0, 1, 2}};
  static const int hex[6][4] = {{0, 1, 5, 4}, {1, 2, 6, 5}, {2, 3, 7, 6}, {3, 0, 4, 7}, {0, 1, 2, 3}, {4, 5, 6, 7}};
  // (dof, order) pairs in the reference's call order: all nodal forces, then all tractions (src/m_global.F90:399-428)
  std::vector<std::pair<int64_t, int32_t>> ent;
  for (int i = 0; i < nfrcs; ++i) {
    ARG(fnode[i] >= 1 && fnode[i] <= c->nnds, "femx_set_loads: force node out of range");
    for (int d = 0; d < pd; ++d) ent.push_back({(int64_t)pd * fnode[i] - pd + d, (i << 3) | d});
  }
  std::vector<int32_t> tel(2 * (size_t)ntrcs, 0);
  for (int i = 0; i < ntrcs; ++i) {
    const int el = telsd[2 * i], side = telsd[2 * i + 1];
    tel[2 * i] = el - 1; tel[2 * i + 1] = side;
    if (el == 0) continue;   // `if (el/=0)`, src/m_global.F90:427
    ARG(el >= 1 && el <= c->nels, "femx_set_loads: traction element out of range");
    const int t = c->etype[el - 1];
    ARG(t != FEMX_COH, "femx_set_loads: traction on a cohesive element");
    const int nsides = t == FEMX_TRI ? 3 : (t == FEMX_HEX ? 6 : 4), nps = t == FEMX_HEX ? 4 : (t == FEMX_TET ? 3 : 2);
    ARG(side >= 1 && side <= nsides, "femx_set_loads: side out of range");
    const int *j = t == FEMX_TRI ? tri[side - 1] : t == FEMX_QUA ? qua[side - 1] : t == FEMX_TET ? tet[side - 1] : hex[side - 1];
    for (int q = 0; q < nps; ++q) {
      const int g = c->nl2g[c->eind[c->eptr[el - 1] + j[q]] - 1];
      for (int d = 0; d < pd; ++d) ent.push_back({(int64_t)pd * g - pd + d, (i << 3) | 4 | d});
    }
  }
  std::stable_sort(ent.begin(), ent.end(), [](const auto &a, const auto &b) { return a.first < b.first; });
  std::vector<int64_t> dof;
  std::vector<int32_t> ptr, src;
  for (size_t k = 0; k < ent.size(); ++k) {
    if (k == 0 || ent[k].first != ent[k - 1].first) { dof.push_back(ent[k].first); ptr.push_back((int32_t)k); }
    src.push_back(ent[k].second);
  }
  ptr.push_back((int32_t)ent.size());
  auto &L = c->loads;
  dev_release(c, L.dof); dev_release(c, L.ptr); dev_release(c, L.src); dev_release(c, L.fval); dev_release(c, L.tval); dev_release(c, L.tel);
  L = femx::DevLoads{};
  L.ntouched = (int)dof.size(); L.nfrcs = nfrcs; L.ntrcs = ntrcs;
  int rc;
  if ((rc = upload(c, &L.dof, dof)) || (rc = upload(c, &L.ptr, ptr)) || (rc = upload(c, &L.src, src))) return rc;
  if ((rc = upload(c, &L.fval, std::vector<double>(fval, fval + (size_t)nfrcs * w)))) return rc;
  if ((rc = upload(c, &L.tval, std::vector<double>(tval, tval + (size_t)ntrcs * w)))) return rc;
  if ((rc = upload(c, &L.tel, tel))) return rc;
  CK(cudaStreamSynchronize(c->stream));
  return FEMX_OK;
}

int femx_form_rhs(femx_ctx *c, double t_init, double dt, double t_total, double *f, int where) {
  if (!c) return FEMX_ERR_ARG;
  ARG(c->has_plan && f, "femx_build_plan first / NULL argument");
  CK(cudaSetDevice(c->device));
  const int64_t nglob = (int64_t)c->pdim * c->nnds;
  int rc;
  double *d_f = nullptr;
  if ((rc = stage_out(c, f, nglob, where, &c->d_vec[3], &d_f))) return rc;
  if ((rc = begin_timing(c))) return rc;
  CK(cudaMemsetAsync(d_f, 0, sizeof(double) * (size_t)nglob, c->stream));   // VecZeroEntries(Vec_F)
  c->launches += launch_form_rhs(c, t_init, dt, d_f);
  c->launches += launch_bc_force(c, dt, t_total, d_f);                      // EnforceBCForce
  if ((rc = end_timing(c))) return rc;
  if ((rc = copy_back(c, f, d_f, nglob, where))) return rc;
  return finish_call(c, true);
}

int femx_gather_sum(femx_ctx *c, const double *a, int64_t na, const double *b, const int64_t *ptr, const int64_t *idx,
                    int64_t nout, double *out) {
  if (!c) return FEMX_ERR_ARG;
  ARG(ptr && idx && out && nout >= 0, "femx_gather_sum: NULL argument");
  CK(cudaSetDevice(c->device));
  int rc;
  if ((rc = begin_timing(c))) return rc;
  c->launches += launch_gather_sum(c, a, na, b, ptr, idx, nout, out);
  if ((rc = end_timing(c))) return rc;
  return finish_call(c, true);
}

static int stress_impl(femx_ctx *c, const double *uu_local, double *stress_node, int layout, int where, int average) {
  ARG(c->has_plan && uu_local && stress_node, "femx_build_plan first / NULL argument");
  CK(cudaSetDevice(c->device));
  const int cp = c->pdim == 2 ? 3 : 6;
  const int64_t nout = (int64_t)c->nlnds * cp;
  int rc;
  const double *d_u = nullptr;
  double *d_out = nullptr;
  if ((rc = stage_in(c, uu_local, (int64_t)c->pdim * c->nlnds, where, &c->d_u, &d_u))) return rc;
  if ((rc = stage_out(c, stress_node, nout, where, &c->d_vec[2], &d_out))) return rc;
  if ((rc = ensure(c, &c->d_sn, (int64_t)c->eptr[c->nels] * cp))) return rc;
  if ((rc = begin_timing(c))) return rc;
  c->launches += launch_mesh_stress(c, d_u, c->d_sn, d_out, layout == FEMX_COL_MAJOR, average);
  if ((rc = end_timing(c))) return rc;
  if ((rc = copy_back(c, stress_node, d_out, nout, where))) return rc;
  return finish_call(c, true);
}

int femx_stress(femx_ctx *c, const double *uu_local, double *stress_node, int layout, int where) {
  if (!c) return FEMX_ERR_ARG;
  return stress_impl(c, uu_local, stress_node, layout, where, 1);
}

int femx_stress_partial(femx_ctx *c, const double *uu_local, double *stress_sum, int32_t *count_node, int where) {
  if (!c) return FEMX_ERR_ARG;
  int rc = stress_impl(c, uu_local, stress_sum, FEMX_ROW_MAJOR, where, 0);
  if (rc || !count_node) return rc;
  CK(cudaMemcpy(count_node, c->dm.node_count, sizeof(int32_t) * (size_t)c->nlnds,
                where == FEMX_DEVICE ? cudaMemcpyDeviceToDevice : cudaMemcpyDeviceToHost));
  return FEMX_OK;
}

}  // extern "C"
