/*
 * oracle.h -- CPU restatement of the defmod hot path.  TEST INFRASTRUCTURE ONLY.
 *
 * Nothing under fem_b200/ may include, link or dlopen this.  Only tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
 * use it, and only as the checker / the timed CPU baseline.
 *
 * Parity status: PINNED against the reference's own fixtures
 *   - examples/tmp.txt:15-33            (Jacobian MatView: pattern of all 16 rows,
 *                                         elastic + penalty + cohesive values)
 *   - examples/SNES_test/elastic/{0,1}_output_000000.vtk (sigma = (105,245,0), analytic)
 *   - examples/CZM_for_instability_test.log:70-112 (UEL cohesive nodal forces)
 * see tests/golden/make_golden.py and tests/test_oracle_golden.py.
 * LAPACK (dgetrf/dgetri) and METIS boundaries are unpinned.
 *
 * Conventions (identical to include/femx.h so tests read the same on both sides):
 *   element type codes = getElTypeNo (src/m_elems.F90:25): 1 tri, 2 qua, 3 tet,
 *   4 hex, 5 coh.  Node ids crossing the API are 1-based LOCAL ids
 *   (src/main.F90:68-73); global dofs are 0-based (src/main.F90:93-94).
 *   coords is row-major (node, dim).  Ke is column-major (Fortran k(:,:)).
 */
#ifndef FEMX_ORACLE_H
#define FEMX_ORACLE_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

enum { ORC_TRI = 1, ORC_QUA = 2, ORC_TET = 3, ORC_HEX = 4, ORC_COH = 5 };

/* quirk switches; 0 = corrected "uel" form, 1 = reference as written */
enum {
  ORC_Q_GATHER_LOCAL14 = 1,  /* D3: gap reads local nodes 1..4 of the rank          */
  ORC_Q_TANGENT_PRECEDENCE = 2, /* D4: lost parentheses in applyStiff_1              */
  ORC_Q_FORCE_LAST_IP = 4,   /* D8: nodal force assigned, only last ip survives     */
  ORC_Q_COH_COUNTS_NODE = 8, /* D5: cohesive elements bump the nodal stress count   */
  ORC_Q_REFERENCE = 15,
  ORC_Q_UEL = 0
};

typedef struct {
  int pdim;            /* 2 or 3 */
  int nels;            /* local elements, ascending global id                       */
  const int *etype;    /* [nels]                                                     */
  const int *eptr;     /* [nels+1] offsets into eind                                 */
  const int *eind;     /* 1-based local node ids                                     */
  const int *emat;     /* [nels] 1-based row of mat, 0 for cohesive                  */
  const int *enlmat;   /* [nels] 0 linear, k>=1 -> cohprops row k                    */
  int nlnds;           /* local nodes                                                */
  const double *coords;/* [nlnds*pdim] row-major                                     */
  const int *nl2g;     /* [nlnds] 1-based global node id                             */
  int nnds;            /* global node count                                          */
  int nmts;
  const double *mat;   /* [nmts*2] (E, nu)                                           */
  int ncoh;
  const double *cohprops; /* [ncoh*6] (sigma_max, dn, dt, q, r, zeta)                */
  int nbcs;
  const int *bcnode;   /* [nbcs*(1+pdim)] global node (1-based), flags (0 = fixed)   */
  const double *bcval; /* [nbcs*pdim]                                                */
} orc_mesh;

/* ---- element library (src/m_elems.F90) ---- */
int orc_nen(int type);
int orc_nip(int type);
int orc_dim(int type);
void orc_sampts(int type, double *pts /*nip*dim*/, double *w /*nip*/);
void orc_shape(int type, int ip /*0-based*/, double *N /*nen*/);
void orc_shape_d(int type, const double *pt, double *dN /*dim*nen row-major (dim,node)*/);
int orc_nodal_stress_inv(int type, double *N2inv /*nen*nip row-major*/);

/* ---- element kernels (src/m_local.F90) ---- */
void orc_dmat(int pdim, double E, double nu, double *D /*cp*cp row-major*/);
void orc_form_el_ke(int type, const double *ecoords /*nen*dim*/, double E, double nu,
                    double *k /*ndof*ndof column-major*/);
void orc_calc_el_stress(int type, const double *ecoords, const double *edisp /*nen*dim*/,
                        double E, double nu, double *estress /*nip*cp row-major*/);
void orc_recover_nodal_stress(int type, const double *estress /*nip*cp*/,
                              double *stress_at_el /*nen*cp row-major*/);

/* ---- cohesive (src/m_seplaw.F90, src/m_elems.F90:658-720, src/m_global.F90:622-845) ---- */
void orc_seplaw_1_tract(const double *props, const double *gap, const double *vgap,
                        double dtime, double *tract);
void orc_seplaw_1_stiff(const double *props, const double *gap, const double *vgap,
                        double dtime, double *stiff /*2x2 row-major (i,j)*/);
void orc_coh_values(const double *ecoords /*4*2*/, double *tangent, double *normal, double *det);
void orc_apply_tract_1(const double *ecoords, const int *nodes /*4, 1-based local*/,
                       const double *u_local, double dt, const double *props, int quirks,
                       double *result /*8*/, double *per_ip /*2*8 or NULL*/);
void orc_apply_stiff_1(const double *ecoords, const int *nodes, const double *u_local,
                       double dt, const double *props, int quirks,
                       double *result /*8*8 row-major (row,col)*/);

/* ---- global assembly with MatSetValues(ADD_VALUES) semantics ---- */
typedef struct orc_csr orc_csr;   /* opaque: growing sorted-row AIJ, then squeezed CSR */
orc_csr *orc_csr_create(int64_t nrows, int prealloc_per_row);
void orc_csr_destroy(orc_csr *);
void orc_csr_add(orc_csr *, int64_t row, int64_t col, double v);      /* MatSetValue ADD    */
void orc_csr_insert(orc_csr *, int64_t row, int64_t col, double v);   /* MatSetValue INSERT */
int64_t orc_csr_nnz(const orc_csr *);
void orc_csr_export(const orc_csr *, int64_t *rowptr, int32_t *colidx, double *vals);
orc_csr *orc_csr_copy(const orc_csr *);
void orc_csr_mult(const orc_csr *, const double *x, double *y);

/* src/main.F90:157-171 (+ApplyKBC src/m_global.F90:264-299 when apply_bc != 0) */
orc_csr *orc_assemble_K(const orc_mesh *, int prealloc_per_row, int apply_bc);
/* src/m_global.F90:98-158: J = copy(K); += cohesive tangents; BC diagonal re-inserted */
orc_csr *orc_jacobian(const orc_mesh *, const orc_csr *K, const double *u_total_local,
                      double dt, int quirks);
/* src/m_global.F90:161-235: R = J*du - (F + sum f_coh, BC rows overwritten) */
void orc_residual(const orc_mesh *, const orc_csr *J, const double *u_total_local,
                  const double *du_global, const double *f_ext_global, double dt,
                  double t_total, int quirks, double *r_global);
/* loads: EdgeAreaNodes* (src/m_elems.F90:282-599), FormRHS + ApplyTraction + EnforceBCForce
 * (src/m_global.F90:343-455).  fval / tval rows are (components[pdim], t1, t2); telsd rows (local element, side). */
int orc_edge_area_nodes(int type, const double *ecoords, int side, double *area, int *slots);
void orc_form_rhs(const orc_mesh *, int nfrcs, const int *fnode, const double *fval, int ntrcs, const int *telsd,
                  const double *tval, double t_init, double dt, double t_total, double *f /*[pdim*nnds]*/);
/* elastic internal force f = sum_e Ke u_e, scattered to global dofs (matrix-free K*u) */
void orc_fint(const orc_mesh *, const double *u_local, double *f_global);
/* src/main.F90:253-279: nodal-averaged stress, [nlnds*cp] row-major */
void orc_stress(const orc_mesh *, const double *uu_local, int quirks, double *stress_node,
                int *count_node);

/* timed CPU baseline: K assembly + f_int, returns seconds; nthreads partitions by element
 * ranges (each thread its own matrix, like one MPI rank per partition). */
double orc_time_assembly(const orc_mesh *, int prealloc_per_row, const double *u_local,
                         int nthreads, int64_t *nnz_out);

#ifdef __cplusplus
}
#endif
#endif
