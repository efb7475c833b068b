// launch.h -- host-callable launchers implemented in the .cu files (each returns the
// number of kernels it launched on ctx->stream).
#pragma once
#include "femx_internal.h"

namespace femx {

void set_n2inv(const double *qua16, const double *hex64);
int launch_elem_rows(femx_ctx *c, const double *d_u, double *d_ke, double *d_fe);
int launch_gather_rows(femx_ctx *c, const double *d_ke, const double *d_fe, double *d_vals, double *d_fint);
int launch_element_ke(femx_ctx *c, int type, int n, const double *d_x, const double *d_enu, double *d_out);
int launch_element_stress(femx_ctx *c, int type, int n, const double *d_x, const double *d_u, const double *d_enu,
                          double *d_sig, double *d_sn);
int launch_mesh_stress(femx_ctx *c, const double *d_uu, double *d_sn, double *d_out, int col_major, int average);
int launch_colidx(femx_ctx *c, int32_t *d_colidx);
int launch_coo(femx_ctx *c, int32_t *d_i, int32_t *d_j);
int build_plan_device(femx_ctx *c);          // plan_gpu.cu: the integer pass on the device (default)
int plan_host_blocks(femx_ctx *c);           // ... host copy of bcol, fetched when an exporter first needs it
int plan_simplex_localize(femx_ctx *c);     // plan_gpu.cu: sx_inc (node ids) -> sx_loc (positions in the row), after either plan pass
int plan_host_bpos(femx_ctx *c);             // ... of bpos
int launch_hex_detj(femx_ctx *c, int *d_bad);   // hex8 meshes: counts (element, integration point) pairs with detJ <= 0
int launch_apply_kbc(femx_ctx *c, double *d_vals);
int launch_bc_force(femx_ctx *c, double dt, double t_total, double *d_out);
int launch_form_rhs(femx_ctx *c, double t_init, double dt, double *d_f);
int launch_spmv_sub(femx_ctx *c, const double *d_vals, const double *d_x, const double *d_sub, double *d_y);
int launch_cohesive_mesh(femx_ctx *c, const double *d_u, double dt, double *d_force, double *d_stiff);
int launch_cohesive_unit(femx_ctx *c, int n, const double *d_x, const int32_t *d_nodes, const double *d_u,
                         const double *d_props, double dt, double *d_force, double *d_stiff);
int launch_jacobian_overlay(femx_ctx *c, const double *d_kvals, double *d_jvals);
int launch_cohesive_force_gather(femx_ctx *c, double *d_out);
int launch_gather_sum(femx_ctx *c, const double *a, int64_t na, const double *b, const int64_t *ptr, const int64_t *idx,
                      int64_t nout, double *out);
int launch_hex_sym6(femx_ctx *c, const double *d_u, double *d_vals, double *d_fint);
int launch_hex_sym7(femx_ctx *c, const double *d_u, double *d_vals, double *d_fint);
int launch_hex_mma(femx_ctx *c, const double *d_u, double *d_vals, double *d_fint);
void hex_mma_grid(const femx_ctx *c, int *gx, int *gy, int *gz, int *nxt, int *nyt, int *lz);
int launch_hex_mma_part(femx_ctx *c, const double *d_u, double *d_vals, double *d_fint, const int *d_list, int n, cudaStream_t st);
int peer_hex_split(femx_ctx *c);            // peer.cu: interface / interior CTA lists of the tile kernel (cached)
int launch_hex_mma_prof(femx_ctx *c, const double *d_u, double *d_vals, double *d_fint);
PeerTab peer_tab(const femx_ctx *c);          // peer.cu
int launch_peer_finish(femx_ctx *c, bool fused, const double *d_vals, double *d_fint);   // push + step barrier + receiver add
bool rows_fused_ok(const femx_ctx *c);
int launch_rows_fused(femx_ctx *c, const double *d_u, double *d_vals, double *d_fint);
bool rows_simplex_ok(const femx_ctx *c);     // rows_simplex.cu
int launch_rows_simplex(femx_ctx *c, const double *d_u, double *d_vals, double *d_fint);

}  // namespace femx
