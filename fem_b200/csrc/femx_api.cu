// femx_api.cu -- the C ABI of include/femx.h.  Plain pointers and sizes in, status codes out.
#include <chrono>

#include "api_util.h"

static std::string g_create_err;


extern "C" {

int femx_version(void) { return 100; }

const char *femx_last_error(const femx_ctx *c) { return c ? c->err.c_str() : g_create_err.c_str(); }

int femx_create(femx_ctx **out, int device, const femx_opts *opts) {
  if (!out) return FEMX_ERR_ARG;
  *out = nullptr;
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || ndev == 0) {
    g_create_err = std::string("no CUDA device: ") + (e != cudaSuccess ? cudaGetErrorString(e) : "device count is 0") +
                   " (femx has no CPU fallback)";
    return FEMX_ERR_CUDA;
  }
  if (device < 0 || device >= ndev) { g_create_err = "device index out of range"; return FEMX_ERR_ARG; }
  if (opts) {
    const int av = opts->assembly_variant;
    bool known = av == 0 || av == 1 || av == 9 || av == 11 || av == 12 || av == 20;
#ifdef FEMX_DIAG
    known = known || av == 13 || av == 21;
#endif
    if (!known) { g_create_err = "unknown assembly_variant (0, 1, 9, 11, 12, 20)"; return FEMX_ERR_ARG; }
  }
  if ((e = cudaSetDevice(device)) != cudaSuccess) { g_create_err = cudaGetErrorString(e); return FEMX_ERR_CUDA; }
  femx_ctx *c = new femx_ctx;
  c->device = device;
  if (opts) c->opts = *opts;
  if (cudaStreamCreateWithFlags(&c->own_stream, cudaStreamNonBlocking) != cudaSuccess || (c->stream = c->own_stream, false) ||
      cudaEventCreate(&c->ev0) != cudaSuccess || cudaEventCreate(&c->ev1) != cudaSuccess) {
    g_create_err = "stream/event creation failed";
    delete c;
    return FEMX_ERR_CUDA;
  }
  double q[16], h[64];
  nodal_stress_tables(q, h);
  set_n2inv(q, h);
  *out = c;
  return FEMX_OK;
}

int femx_destroy(femx_ctx *c) {
  if (!c) return FEMX_OK;
  cudaSetDevice(c->device);
  cudaStreamSynchronize(c->stream);
  femx_peer_disconnect(c);
  if (c->peer.arena) cudaFree(c->peer.arena);
  for (void *p : c->allocs) cudaFree(p);
  cudaEventDestroy(c->ev0);
  cudaEventDestroy(c->ev1);
  if (c->ev_mid) cudaEventDestroy(c->ev_mid);
  if (c->peer.hi) { cudaStreamDestroy(c->peer.hi); cudaEventDestroy(c->peer.ev_fork); cudaEventDestroy(c->peer.ev_join); }
  cudaStreamDestroy(c->own_stream);
  delete c;
  return FEMX_OK;
}

int femx_host_alloc(void **ptr, int64_t bytes) {
  return cudaMallocHost(ptr, (size_t)(bytes > 0 ? bytes : 1)) == cudaSuccess ? FEMX_OK : FEMX_ERR_ALLOC;
}
int femx_host_free(void *ptr) { return cudaFreeHost(ptr) == cudaSuccess ? FEMX_OK : FEMX_ERR_CUDA; }

int femx_device_alloc(femx_ctx *c, void **ptr, int64_t bytes) {
  if (!c || !ptr) return FEMX_ERR_ARG;
  char *p = nullptr;
  int rc = dev_alloc(c, &p, bytes);
  *ptr = p;
  return rc;
}
int femx_device_free(femx_ctx *c, void *ptr) {
  if (!c) return FEMX_ERR_ARG;
  dev_release(c, ptr);
  return FEMX_OK;
}
int femx_memcpy(femx_ctx *c, void *dst, const void *src, int64_t bytes, int dw, int sw) {
  if (!c) return FEMX_ERR_ARG;
  cudaMemcpyKind k = dw == FEMX_DEVICE ? (sw == FEMX_DEVICE ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice)
                                       : (sw == FEMX_DEVICE ? cudaMemcpyDeviceToHost : cudaMemcpyHostToHost);
  CK(cudaMemcpyAsync(dst, src, (size_t)bytes, k, c->stream));
  CK(cudaStreamSynchronize(c->stream));
  return FEMX_OK;
}
int femx_synchronize(femx_ctx *c) {
  if (!c) return FEMX_ERR_ARG;
  CK(cudaStreamSynchronize(c->stream));
  CK(cudaGetLastError());
  if (c->ms_pending) {
    float ms = 0.f;
    if (cudaEventElapsedTime(&ms, c->ev0, c->ev1) == cudaSuccess) c->last_ms = ms;
    c->ms_pending = false;
  }
  if (c->peer.connected && c->peer.d_err && !c->peer.deferred) return femx_peer_last_error(c);
  return FEMX_OK;
}
// Threading / streams: by default every call runs on the ctx's own stream and returns synchronised.
// femx_set_stream puts the ctx on the caller's CUDA stream (a cudaStream_t passed as void *; NULL = back to its own), so
// that femx work is ordered with the caller's own kernels / PETSc's; femx_set_async(1) makes calls on FEMX_DEVICE buffers
// return as soon as the work is enqueued -- the host overlaps femx_jacobian with anything else and calls femx_synchronize
// (or synchronises its stream) before it reads results or timings.  Host-buffer calls stay usable: their D2H copies are
// enqueued too and complete at femx_synchronize.
int femx_set_stream(femx_ctx *c, void *cuda_stream) {
  if (!c) return FEMX_ERR_ARG;
  CK(cudaStreamSynchronize(c->stream));
  c->stream = cuda_stream ? (cudaStream_t)cuda_stream : c->own_stream;
  return FEMX_OK;
}
int femx_set_async(femx_ctx *c, int on) {
  if (!c) return FEMX_ERR_ARG;
  if (!on) CK(cudaStreamSynchronize(c->stream));
  c->async = on != 0;
  return FEMX_OK;
}
int64_t femx_launch_count(const femx_ctx *c) { return c ? c->launches : 0; }
double femx_last_kernel_ms(const femx_ctx *c) { return c ? c->last_ms : 0.0; }
double femx_plan_seconds(const femx_ctx *c) { return c ? c->plan_seconds : 0.0; }
double femx_last_assembly_kernel_ms(femx_ctx *c) {
  if (!c) return 0.0;
  if (!c->mid_valid) return c->last_ms;     // no peer exchange behind the kernel: the whole timed region is the kernel
  float ms = 0.f;
  if (cudaEventSynchronize(c->ev_mid) != cudaSuccess || cudaEventElapsedTime(&ms, c->ev0, c->ev_mid) != cudaSuccess) { cudaGetLastError(); return 0.0; }
  return ms;
}

// --------------------------------------------------------------------------- state
int femx_set_mesh(femx_ctx *c, int pdim, int nels, const int *etype, const int *eptr, const int *eind,
                  const int *emat, const int *enlmat, int nlnds, const double *coords, int layout,
                  const int *nl2g, int nnds) {
  if (!c) return FEMX_ERR_ARG;
  ARG(pdim == 2 || pdim == 3, "pdim must be 2 or 3");
  ARG(nels >= 0 && nlnds >= 0 && nnds >= nlnds, "bad sizes");
  ARG(etype && eptr && eind && emat && enlmat && coords && nl2g, "null mesh array");
  // validate everything before the ctx is touched: a refused mesh leaves the previous mesh and plan intact
  ARG(eptr[0] == 0, "eptr must start at 0");
  for (int e = 0; e < nels; ++e) {
    ARG(etype[e] >= FEMX_TRI && etype[e] <= FEMX_COH, "unknown element type");
    ARG(eptr[e + 1] - eptr[e] == femx::type_nen(etype[e]), "eptr does not match the element types' node counts");
    ARG(type_dim(etype[e]) == pdim, "element type does not match pdim");
    for (int q = eptr[e]; q < eptr[e + 1]; ++q) ARG(eind[q] >= 1 && eind[q] <= nlnds, "element node id out of range (1-based local ids)");
  }
  for (int i = 0; i < nlnds; ++i) {
    ARG(nl2g[i] >= 1 && nl2g[i] <= nnds, "nl2g out of range (1-based global ids)");
    ARG(i == 0 || nl2g[i] > nl2g[i - 1], "nl2g must ascend (src/main.F90:57-63 numbering)");
  }
  c->has_plan = false;
  c->pdim = pdim; c->nels = nels; c->nlnds = nlnds; c->nnds = nnds;
  c->etype.assign(etype, etype + nels);
  c->eptr.assign(eptr, eptr + nels + 1);
  c->eind.assign(eind, eind + eptr[nels]);
  c->emat.assign(emat, emat + nels);
  c->enlmat.assign(enlmat, enlmat + nels);
  c->nl2g.assign(nl2g, nl2g + nlnds);
  c->coords.resize((size_t)nlnds * pdim);
  for (int i = 0; i < nlnds; ++i)
    for (int d = 0; d < pdim; ++d)
      c->coords[(size_t)i * pdim + d] = layout == FEMX_COL_MAJOR ? coords[(size_t)d * nlnds + i] : coords[(size_t)i * pdim + d];
  c->has_mesh = true;
  c->has_plan = false;
  return FEMX_OK;
}

// Everything the kernels derive from (E, nu): (lambda, mu) per material, and for a logical hex8 grid sqrt(mu) per cell and
// rho = lambda / mu.  Runs in femx_build_plan and again when femx_set_materials is called afterwards, so that the fast
// kernels never see scale factors of an older material table (they are re-validated too: one Poisson ratio, mu > 0).
static int refresh_materials(femx_ctx *c) {
  for (int e = 0; e < c->nels; ++e) {
    if (c->etype[e] != FEMX_COH) ARG(c->emat[e] >= 1 && c->emat[e] <= c->nmts, "element material id out of range (femx_set_materials first)");
    if (c->enlmat[e] != 0) ARG(c->enlmat[e] >= 1 && c->enlmat[e] <= c->ncoh, "nlMat out of range");
  }
  int rc;
  std::vector<double> lm(2 * (size_t)c->nmts);
  for (int m = 0; m < c->nmts; ++m) lame(c->E_nu[2 * m], c->E_nu[2 * m + 1], &lm[2 * m], &lm[2 * m + 1]);
  dev_release(c, c->dm.lam_mu); dev_release(c, c->dm.cohprops);
  c->dm.lam_mu = nullptr; c->dm.cohprops = nullptr;
  if ((rc = upload(c, &c->dm.lam_mu, lm))) return rc;
  if ((rc = upload(c, &c->dm.cohprops, c->cohprops))) return rc;
  std::vector<char> used(c->nmts, 0);
  for (int e = 0; e < c->nels; ++e)
    if (c->etype[e] != FEMX_COH) used[c->emat[e] - 1] = 1;
  // both fast paths scale gradients by sqrt(mu): every used material needs mu > 0 and a finite lambda / mu
  bool mu_ok = true;
  for (int m = 0; m < c->nmts; ++m)
    if (used[m] && (!(lm[2 * m + 1] > 0.0) || !std::isfinite(lm[2 * m] / lm[2 * m + 1]))) mu_ok = false;
  c->mat_ok_simplex = mu_ok;
  c->mat_ok_hex = false;
  c->mat_why.clear();
  if (c->hex.ok) {
    femx::DevHex &hx = c->dhex;
    double nu0 = 0.0;
    bool have = false, one_nu = true;
    for (int m = 0; m < c->nmts; ++m)
      if (used[m]) {
        if (!have) { nu0 = c->E_nu[2 * m + 1]; have = true; }
        else if (c->E_nu[2 * m + 1] != nu0) one_nu = false;
      }
    if (!one_nu) c->mat_why = "elements with different Poisson ratios";
    else if (!mu_ok) c->mat_why = "a material with mu <= 0";
    else {
      c->mat_ok_hex = true;
      hx.rho = 2.0 * nu0 / (1.0 - 2.0 * nu0);   // lambda / mu
      std::vector<double> smu(c->nmts);
      for (int m = 0; m < c->nmts; ++m) smu[m] = std::sqrt(lm[2 * m + 1] > 0.0 ? lm[2 * m + 1] : 0.0);
      std::vector<double> cm(c->hex.elem_at.size(), 0.0);
      for (size_t q = 0; q < cm.size(); ++q)
        if (c->hex.elem_at[q] >= 0) cm[q] = smu[c->emat[c->hex.elem_at[q]] - 1];
      dev_release(c, hx.sqrt_mu); dev_release(c, hx.cellmu);
      hx.sqrt_mu = nullptr; hx.cellmu = nullptr;
      if ((rc = upload(c, &hx.sqrt_mu, smu))) return rc;
      if ((rc = upload(c, &hx.cellmu, cm))) return rc;
    }
  }
  CK(cudaStreamSynchronize(c->stream));
  return FEMX_OK;
}

int femx_set_materials(femx_ctx *c, int nmts, const double *E_nu, int ncoh, const double *cohprops) {
  if (!c) return FEMX_ERR_ARG;
  ARG(nmts >= 0 && ncoh >= 0, "bad counts");
  c->nmts = nmts; c->ncoh = ncoh;
  c->E_nu.assign(E_nu, E_nu + 2 * (size_t)nmts);
  c->cohprops.assign(cohprops, cohprops + 6 * (size_t)ncoh);
  if (c->has_plan) {   // a material change after the plan: re-validate and recompute every derived factor
    int rc = refresh_materials(c);
    if (rc) { c->has_plan = false; return rc; }
  }
  return FEMX_OK;
}

static int prepare_bcs(femx_ctx *c) {
  dev_release(c, c->d_bc_diag); dev_release(c, c->d_bc_dof); dev_release(c, c->d_bc_val);
  c->d_bc_diag = nullptr; c->d_bc_dof = nullptr; c->d_bc_val = nullptr;
  std::vector<int64_t> diag, dof;
  std::vector<double> val;
  const int pd = c->pdim;
  for (int i = 0; i < c->nbcs; ++i) {
    const int node = c->bcnode[(size_t)i * (1 + pd)];
    auto it = std::lower_bound(c->nl2g.begin(), c->nl2g.end(), node);
    if (it == c->nl2g.end() || *it != node) continue;  // node not on this rank
    const int r = (int)(it - c->nl2g.begin());
    const int64_t b0 = c->plan.bptr[r];
    const int nb = (int)(c->plan.bptr[r + 1] - b0);
    for (int d = 0; d < pd; ++d)
      if (c->bcnode[(size_t)i * (1 + pd) + 1 + d] == 0) {  // BC_PRESENT
        diag.push_back((int64_t)pd * pd * b0 + (int64_t)d * pd * nb + pd * c->plan.bdiag[r] + d);
        dof.push_back((int64_t)pd * node - pd + d);
        val.push_back(c->bcval[(size_t)i * pd + d]);
      }
  }
  c->nbc_dofs = (int)diag.size();
  int rc;
  if ((rc = upload(c, &c->d_bc_diag, diag))) return rc;
  if ((rc = upload(c, &c->d_bc_dof, dof))) return rc;
  if ((rc = upload(c, &c->d_bc_val, val))) return rc;
  CK(cudaStreamSynchronize(c->stream));
  return FEMX_OK;
}

int femx_set_bcs(femx_ctx *c, int nbcs, const int *bcnode, const double *bcval) {
  if (!c) return FEMX_ERR_ARG;
  ARG(c->has_mesh, "femx_set_mesh first");
  ARG(nbcs >= 0, "bad nbcs");
  c->nbcs = nbcs;
  c->bcnode.assign(bcnode, bcnode + (size_t)nbcs * (1 + c->pdim));
  c->bcval.assign(bcval, bcval + (size_t)nbcs * c->pdim);
  return c->has_plan ? prepare_bcs(c) : FEMX_OK;
}

int femx_set_ownership(femx_ctx *c, int nranks, int rank, int64_t row_start, int64_t row_end) {
  if (!c) return FEMX_ERR_ARG;
  ARG(nranks >= 1 && rank >= 0 && rank < nranks && row_start >= 0 && row_end >= row_start, "bad ownership");
  c->own_nranks = nranks; c->own_rank = rank; c->own_row_start = row_start; c->own_row_end = row_end;
  return FEMX_OK;
}
// --------------------------------------------------------------------------- plan
int femx_build_plan(femx_ctx *c) {
  if (!c) return FEMX_ERR_ARG;
  ARG(c->has_mesh, "femx_set_mesh first");
  CK(cudaSetDevice(c->device));
  const auto t_plan0 = std::chrono::steady_clock::now();
  // a second plan on the same ctx: the previous plan's device arrays go first (hundreds of MB at 1 M hex8)
  c->has_plan = false;
  CK(cudaStreamSynchronize(c->stream));
  for (void *q : c->plan_allocs) dev_release(c, q);
  c->plan_allocs.clear();
  for (double **q : {&c->d_ke, &c->d_fe, &c->d_sn, &c->d_u, &c->d_vals, &c->d_kstage, &c->d_vec[0], &c->d_vec[1], &c->d_vec[2], &c->d_vec[3]}) {
    dev_release(c, *q);
    *q = nullptr;
  }
  c->buf_n.clear();
  c->dm = femx::DevMesh(); c->dhex = femx::DevHex(); c->loads = femx::DevLoads();
  c->d_bc_diag = nullptr; c->d_bc_dof = nullptr; c->d_bc_val = nullptr;
  const std::vector<void *> before = c->allocs;
  int rc;
  c->peer.split_valid = false;
  const bool host_plan = c->opts.reserved[4] == 1;   // plan.cpp / plan_hex.cpp on the host: the cross-check of the device pass (plan_gpu.cu)
  if (!host_plan) {
    if ((rc = build_plan_device(c))) return rc;
    if ((rc = plan_simplex_localize(c))) return rc;
    if (c->hex.ok) {
      // the tile kernels scale gradients by rsqrt(detJ): a mirrored grid or an inverted element (which the reference and the
      // general kernels assemble as is, src/m_local.F90:61-65) must not reach them
      int *d_bad = nullptr, bad = 0;
      if ((rc = dev_alloc(c, &d_bad, 1))) return rc;
      CK(cudaMemsetAsync(d_bad, 0, sizeof(int), c->stream));
      c->launches += launch_hex_detj(c, d_bad);
      CK(cudaMemcpyAsync(&bad, d_bad, sizeof(int), cudaMemcpyDeviceToHost, c->stream));
      CK(cudaStreamSynchronize(c->stream));
      dev_release(c, d_bad);
      if (bad) { c->hex.ok = false; c->hex.why = "an element with detJ <= 0 (inverted or mirrored numbering)"; }
    }
    const int av = c->opts.assembly_variant;
    if (!c->hex.ok && (av == 12 || av == 13 || av == 20 || av == 21)) {
      c->err = "fused assembly variant requested but the mesh does not qualify: " + c->hex.why;
      return FEMX_ERR_UNSUPPORTED;
    }
  } else {
  c->host_bcol = true; c->host_bpos = true;
  rc = build_plan(c->pdim, c->nels, c->etype.data(), c->eptr.data(), c->eind.data(), c->enlmat.data(),
                      c->nlnds, c->nnds, &c->plan, &c->err);
  if (rc) return rc;
  for (int e = 0; e < c->nels; ++e)
    if (c->etype[e] != FEMX_COH) ARG(c->emat[e] >= 1 && c->emat[e] <= c->nmts, "element material id out of range (femx_set_materials first)");
  femx::Plan &P = c->plan;
  c->nl2g_identity = c->nlnds == c->nnds;
  for (int i = 0; i < c->nlnds && c->nl2g_identity; ++i) c->nl2g_identity = c->nl2g[i] == i + 1;
  if (P.simplex) {   // the simplex rows kernel scales gradients by sqrt(w detJ mu): admit only positive Jacobians (mu > 0: refresh_materials)
    const int pd = c->pdim;
    for (int e = 0; e < c->nels && P.simplex; ++e) {
      const int *nd = &c->eind[c->eptr[e]];
      double d[3][3];
      for (int b = 1; b <= pd; ++b)
        for (int k = 0; k < pd; ++k) d[b - 1][k] = c->coords[(size_t)(nd[b] - 1) * pd + k] - c->coords[(size_t)(nd[0] - 1) * pd + k];
      const double det = pd == 2 ? d[0][0] * d[1][1] - d[0][1] * d[1][0]
                                 : d[0][0] * (d[1][1] * d[2][2] - d[1][2] * d[2][1]) - d[0][1] * (d[1][0] * d[2][2] - d[1][2] * d[2][0]) +
                                       d[0][2] * (d[1][0] * d[2][1] - d[1][1] * d[2][0]);
      if (!(det > 0.0)) P.simplex = false;
    }
    if (!P.simplex) { P.sx_ent.clear(); P.sx_inc.clear(); P.sx_mat.clear(); }
  }
  for (size_t p = 0; p < P.sx_mat.size(); ++p) P.sx_mat[p] = c->emat[P.n2e_elem[p]] - 1;
  P.ri_nodes.clear(); P.ri_info.clear();
  if (c->pdim == 2 && P.max_inc >= 1 && P.max_inc <= 32) {   // flattened incidence records of the fused rows kernel
    const size_t ninc = (size_t)P.n2e_ptr[c->nlnds];
    P.ri_nodes.assign(4 * ninc, 0); P.ri_info.assign(ninc, 0);
    for (size_t p = 0; p < ninc; ++p) {
      const int e = P.n2e_elem[p], nen = c->eptr[e + 1] - c->eptr[e];
      for (int b = 0; b < nen; ++b) P.ri_nodes[4 * p + b] = P.conn0[c->eptr[e] + b];
      const int m = c->etype[e] == FEMX_COH ? 0 : c->emat[e] - 1;
      P.ri_info[p] = c->etype[e] | (P.n2e_a[p] << 4) | (m << 8);
    }
  }
  femx::DevMesh &dm = c->dm;
  dm.pdim = c->pdim; dm.nels = c->nels; dm.nlnds = c->nlnds; dm.nnds = c->nnds; dm.cp = c->pdim == 2 ? 3 : 6;
  c->nnz = (int64_t)c->pdim * c->pdim * P.nblk;
#define UP(dst, src) if ((rc = upload(c, &dst, src))) return rc
  UP(dm.etype, c->etype); UP(dm.eptr, c->eptr); UP(dm.conn0, P.conn0); UP(dm.emat, c->emat); UP(dm.enlmat, c->enlmat);
  UP(dm.nl2g, c->nl2g); UP(dm.coords, c->coords);
  UP(dm.bptr, P.bptr); UP(dm.bcol, P.bcol); UP(dm.bdiag, P.bdiag); UP(dm.n2e_ptr, P.n2e_ptr); UP(dm.n2e_elem, P.n2e_elem);
  UP(dm.rbe_ptr, P.rbe_ptr); UP(dm.rb_off, P.rb_off); UP(dm.rb_ent, P.rb_ent); UP(dm.sx_ent, P.sx_ent); UP(dm.sx_inc, P.sx_inc); UP(dm.sx_mat, P.sx_mat); UP(dm.ri_nodes, P.ri_nodes); UP(dm.ri_info, P.ri_info);
  UP(dm.n2e_a, P.n2e_a); UP(dm.eb_ptr, P.eb_ptr); UP(dm.bpos, P.bpos); UP(dm.ke_off, P.ke_off); UP(dm.fe_off, P.fe_off);
  if ((rc = plan_simplex_localize(c))) return rc;
  // stress multiplicity (count_node, src/main.F90:263); mesh constant, computed once
  std::vector<int32_t> cnt(c->nlnds, 0);
  for (int r = 0; r < c->nlnds; ++r)
    for (int64_t p = P.n2e_ptr[r]; p < P.n2e_ptr[r + 1]; ++p)
      if (c->etype[P.n2e_elem[p]] != FEMX_COH || (c->opts.quirks & FEMX_Q_COH_COUNTS_NODE)) cnt[r]++;
  UP(dm.node_count, cnt);
  // type buckets
  c->buckets.clear();
  for (int t = FEMX_TRI; t <= FEMX_COH; ++t) {
    femx::Bucket b;
    b.type = t;
    for (int e = 0; e < c->nels; ++e)
      if (c->etype[e] == t) b.h_elem.push_back(e);
    b.n = (int)b.h_elem.size();
    if (b.n) { UP(b.d_elem, b.h_elem); c->buckets.push_back(b); }
  }
  UP(c->d_coh_elems, P.coh_elems); UP(c->d_cblk_id, P.cblk_id); UP(c->d_cblk_row, P.cblk_row);
  UP(c->d_cblk_ptr, P.cblk_ptr); UP(c->d_cblk_src, P.cblk_src); UP(c->d_cnode_id, P.cnode_id);
  UP(c->d_cnode_ptr, P.cnode_ptr); UP(c->d_cnode_src, P.cnode_src);
  if ((rc = dev_alloc(c, &c->d_coh_ke, 64 * (int64_t)P.coh_elems.size()))) return rc;
  if ((rc = dev_alloc(c, &c->d_coh_fe, 8 * (int64_t)P.coh_elems.size()))) return rc;
  // fused hex8 tile kernels (variants 12, 20): only if the mesh provably is a logical grid of hex8
  c->hex = femx::HexPlan();
  if (c->opts.assembly_variant != 1 && c->opts.assembly_variant != 9 && c->opts.assembly_variant != 11) {
    build_hex_plan(P, c->etype.data(), c->eptr.data(), &c->hex, &c->hex.why);
    if (c->hex.ok) {
      // the tile kernels scale gradients by rsqrt(detJ): a mirrored grid or an inverted element (which the reference and the
      // general kernels assemble as is, src/m_local.F90:61-65) must not reach them
      int *d_bad = nullptr, bad = 0;
      if ((rc = dev_alloc(c, &d_bad, 1))) return rc;
      CK(cudaMemsetAsync(d_bad, 0, sizeof(int), c->stream));
      c->launches += launch_hex_detj(c, d_bad);
      CK(cudaMemcpyAsync(&bad, d_bad, sizeof(int), cudaMemcpyDeviceToHost, c->stream));
      CK(cudaStreamSynchronize(c->stream));
      dev_release(c, d_bad);
      if (bad) { c->hex.ok = false; c->hex.why = "an element with detJ <= 0 (inverted or mirrored numbering)"; }
    }
    if (c->hex.ok) {
      femx::DevHex &hx = c->dhex;
      hx.NI = c->hex.NI; hx.NJ = c->hex.NJ; hx.NK = c->hex.NK;
      UP(hx.node_at, c->hex.node_at); UP(hx.elem_at, c->hex.elem_at); UP(hx.kmap, c->hex.kmap);
      std::vector<double> c4((size_t)c->nlnds * 4, 0.0);
      for (int i = 0; i < c->nlnds; ++i)
        for (int d = 0; d < 3; ++d) c4[(size_t)i * 4 + d] = c->coords[(size_t)i * 3 + d];
      UP(hx.coords4, c4);
      std::vector<double> px(c->hex.node_at.size() * 3, 0.0);
      std::vector<femx::HexRowMeta> pm(c->hex.node_at.size());
      for (size_t q = 0; q < pm.size(); ++q) {
        const int r = c->hex.node_at[q];
        femx::HexRowMeta &mm = pm[q];
        mm.base = 0; mm.nb = 0; mm.grow = 0;
        for (int t = 0; t < 28; ++t) mm.k[t] = 255;
        if (r < 0) continue;
        for (int d = 0; d < 3; ++d) px[q * 3 + d] = c->coords[(size_t)r * 3 + d];
        mm.base = 9 * P.bptr[r];
        mm.nb = (int)(P.bptr[r + 1] - P.bptr[r]);
        mm.grow = 3 * (c->nl2g[r] - 1);
        for (int t = 0; t < 27; ++t) mm.k[t] = c->hex.kmap[(size_t)r * 32 + t];
      }
      UP(hx.pos_x, px); UP(hx.pos_meta, pm);
    } else if (c->opts.assembly_variant == 12 || c->opts.assembly_variant == 13 || c->opts.assembly_variant == 20 || c->opts.assembly_variant == 21) {
      c->err = "fused assembly variant requested but the mesh does not qualify: " + c->hex.why;
      return FEMX_ERR_UNSUPPORTED;
    }
  }
#undef UP
  }
  if ((rc = refresh_materials(c))) return rc;
  const int av = c->opts.assembly_variant;
  if ((av == 12 || av == 13 || av == 20 || av == 21) && !c->mat_ok_hex) {
    c->err = "fused assembly variant requested but the materials do not qualify: " + c->mat_why;
    return FEMX_ERR_UNSUPPORTED;
  }
  CK(cudaStreamSynchronize(c->stream));
  rc = prepare_bcs(c);
  for (void *q : c->allocs)
    if (std::find(before.begin(), before.end(), q) == before.end()) c->plan_allocs.push_back(q);
  if (rc) return rc;
  c->has_plan = true;
  c->plan_seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t_plan0).count();
  return FEMX_OK;
}

int femx_pattern_size(femx_ctx *c, int64_t *nrows, int64_t *nnz) {
  if (!c) return FEMX_ERR_ARG;
  ARG(c->has_plan, "femx_build_plan first");
  if (nrows) *nrows = (int64_t)c->pdim * c->nnds;
  if (nnz) *nnz = c->nnz;
  return FEMX_OK;
}

int femx_get_pattern(femx_ctx *c, int64_t *rowptr, int32_t *colidx, int where) {
  if (!c) return FEMX_ERR_ARG;
  ARG(c->has_plan, "femx_build_plan first");
  const int pd = c->pdim;
  const int64_t n = (int64_t)pd * c->nnds;
  if (rowptr) {
    std::vector<int64_t> rp(n + 1, 0);
    for (int r = 0; r < c->nlnds; ++r) {
      const int64_t nb = c->plan.bptr[r + 1] - c->plan.bptr[r];
      for (int i = 0; i < pd; ++i) rp[(int64_t)pd * (c->nl2g[r] - 1) + i + 1] = pd * nb;
    }
    for (int64_t i = 0; i < n; ++i) rp[i + 1] += rp[i];
    if (where == FEMX_DEVICE) CK(cudaMemcpy(rowptr, rp.data(), sizeof(int64_t) * (n + 1), cudaMemcpyHostToDevice));
    else memcpy(rowptr, rp.data(), sizeof(int64_t) * (n + 1));
  }
  if (colidx) {
    View v(c, colidx, sizeof(int32_t) * (size_t)c->nnz, where, false, true);
    if (v.rc) return v.rc;
    c->launches += launch_colidx(c, v.as<int32_t>());
    int rc = v.finish();
    if (rc) return rc;
    CK(cudaStreamSynchronize(c->stream));
    CK(cudaGetLastError());
  }
  return FEMX_OK;
}

// ---- PETSc hand-off: the index arrays of MatSetPreallocationCOO and the MPIAIJ split of the owned rows
int femx_get_coo(femx_ctx *c, int32_t *coo_i, int32_t *coo_j, int where) {
  if (!c) return FEMX_ERR_ARG;
  ARG(c->has_plan && coo_i && coo_j, "femx_build_plan first / NULL argument");
  ARG((int64_t)c->pdim * c->nnds < 2147483647ll, "global dofs exceed 32-bit PetscInt");
  CK(cudaSetDevice(c->device));
  View vi(c, coo_i, sizeof(int32_t) * (size_t)c->nnz, where, false, true), vj(c, coo_j, sizeof(int32_t) * (size_t)c->nnz, where, false, true);
  if (vi.rc || vj.rc) return vi.rc | vj.rc;
  c->launches += launch_coo(c, vi.as<int32_t>(), vj.as<int32_t>());
  int rc;
  if ((rc = vi.finish()) || (rc = vj.finish())) return rc;
  CK(cudaStreamSynchronize(c->stream));
  CK(cudaGetLastError());
  return FEMX_OK;
}

// rows [row_start, row_end) of the rank-local pattern in MPIAIJ form: the diagonal block (columns inside the range, local
// column numbers) and the off-diagonal block compacted through garray (ascending global columns) -- Mat_MPIAIJ's A / B / garray
static int mpiaij_walk(femx_ctx *c, int64_t row_start, int64_t row_end, int64_t *d_rowptr, int32_t *d_col, int64_t *o_rowptr, int32_t *o_col,
                       std::vector<int32_t> *garray, int64_t *nd_out, int64_t *no_out) {
  const int pd = c->pdim;
  ARG(row_start >= 0 && row_end >= row_start && row_end <= (int64_t)pd * c->nnds, "bad row range");
  { int rcb = plan_host_blocks(c); if (rcb) return rcb; }
  const femx::Plan &P = c->plan;
  std::vector<int32_t> &ga = *garray;
  const bool fill = d_col != nullptr;
  if (!fill) {   // first pass: collect the off-diagonal columns
    ga.clear();
    for (int r = 0; r < c->nlnds; ++r) {
      const int64_t g0 = (int64_t)pd * (c->nl2g[r] - 1);
      if (g0 + pd <= row_start || g0 >= row_end) continue;
      for (int64_t b = P.bptr[r]; b < P.bptr[r + 1]; ++b) {
        const int64_t c0 = (int64_t)pd * (c->nl2g[P.bcol[b]] - 1);
        for (int j = 0; j < pd; ++j)
          if (c0 + j < row_start || c0 + j >= row_end) ga.push_back((int32_t)(c0 + j));
      }
    }
    std::sort(ga.begin(), ga.end());
    ga.erase(std::unique(ga.begin(), ga.end()), ga.end());
  }
  int64_t nd = 0, no = 0;
  int r = (int)(std::lower_bound(c->nl2g.begin(), c->nl2g.end(), (int32_t)(row_start / pd + 1)) - c->nl2g.begin());
  for (int64_t g = row_start; g < row_end; ++g) {
    if (fill) { d_rowptr[g - row_start] = nd; o_rowptr[g - row_start] = no; }
    const int32_t node = (int32_t)(g / pd + 1);
    while (r < c->nlnds && c->nl2g[r] < node) ++r;
    if (r >= c->nlnds || c->nl2g[r] != node) continue;      // a row of a node this rank does not hold: empty
    for (int64_t b = P.bptr[r]; b < P.bptr[r + 1]; ++b) {
      const int64_t c0 = (int64_t)pd * (c->nl2g[P.bcol[b]] - 1);
      for (int j = 0; j < pd; ++j) {
        const int64_t col = c0 + j;
        if (col >= row_start && col < row_end) { if (fill) d_col[nd] = (int32_t)(col - row_start); ++nd; }
        else { if (fill) o_col[no] = (int32_t)(std::lower_bound(ga.begin(), ga.end(), (int32_t)col) - ga.begin()); ++no; }
      }
    }
  }
  if (fill) { d_rowptr[row_end - row_start] = nd; o_rowptr[row_end - row_start] = no; }
  if (nd_out) *nd_out = nd;
  if (no_out) *no_out = no;
  return FEMX_OK;
}

int femx_mpiaij_sizes(femx_ctx *c, int64_t row_start, int64_t row_end, int64_t *nnz_diag, int64_t *nnz_offdiag, int64_t *n_garray) {
  if (!c) return FEMX_ERR_ARG;
  ARG(c->has_plan, "femx_build_plan first");
  std::vector<int32_t> ga;
  int rc = mpiaij_walk(c, row_start, row_end, nullptr, nullptr, nullptr, nullptr, &ga, nnz_diag, nnz_offdiag);
  if (n_garray) *n_garray = (int64_t)ga.size();
  return rc;
}

int femx_get_pattern_mpiaij(femx_ctx *c, int64_t row_start, int64_t row_end, int64_t *d_rowptr, int32_t *d_col, int64_t *o_rowptr, int32_t *o_col,
                            int32_t *garray) {
  if (!c) return FEMX_ERR_ARG;
  ARG(c->has_plan && d_rowptr && d_col && o_rowptr && o_col && garray, "femx_build_plan first / NULL argument");
  std::vector<int32_t> ga;
  int rc = mpiaij_walk(c, row_start, row_end, nullptr, nullptr, nullptr, nullptr, &ga, nullptr, nullptr);
  if (rc) return rc;
  memcpy(garray, ga.data(), sizeof(int32_t) * ga.size());
  return mpiaij_walk(c, row_start, row_end, d_rowptr, d_col, o_rowptr, o_col, &ga, nullptr, nullptr);
}

int femx_get_slot_map(femx_ctx *c, int64_t *offsets, uint16_t *blockpos) {
  if (!c) return FEMX_ERR_ARG;
  ARG(c->has_plan, "femx_build_plan first");
  { int rcb = plan_host_bpos(c); if (rcb) return rcb; }
  if (offsets) memcpy(offsets, c->plan.eb_ptr.data(), sizeof(int64_t) * (c->nels + 1));
  if (blockpos) memcpy(blockpos, c->plan.bpos.data(), sizeof(uint16_t) * c->plan.bpos.size());
  return FEMX_OK;
}

}  // extern "C"
