/* metis_part.c -- the reference's partitioner call, for meshes that go through femx's multi-rank path.
 *
 * PartitionBroadcast (src/m_io.F90:125-141) hands the element -> node lists (0-based) to METIS_PartMeshNodal with every
 * optional argument null and keeps epart.  This wrapper makes the same call on the METIS that ships inside the CUDA
 * toolkit (libmetis_static.a; tools/metis_probe.c shows its idx_t is 64-bit, so the int32 arrays are widened first).
 * Host-side tooling only: nothing under fem_b200/csrc links it.
 *   make -C tools metis   ->  tools/libmetispart.so
 */
#include <stdint.h>
#include <stdlib.h>

int METIS_PartMeshNodal(int64_t *ne, int64_t *nn, int64_t *eptr, int64_t *eind, int64_t *vwgt, int64_t *vsize, int64_t *nparts,
                        float *tpwgts, int64_t *options, int64_t *objval, int64_t *epart, int64_t *npart);

/* returns the METIS status (1 = METIS_OK), -1 on allocation failure; epart[ne] gets the 0-based part of every element */
int femx_metis_part_mesh_nodal(int ne, int nn, const int *eptr, const int *eind, int nparts, int *epart, long long *objval_out) {
  int64_t ne64 = ne, nn64 = nn, np64 = nparts, objval = 0;
  if (nparts <= 1) {
    for (int e = 0; e < ne; ++e) epart[e] = 0;
    if (objval_out) *objval_out = 0;
    return 1;
  }
  const int64_t nind = eptr[ne];
  int64_t *p = malloc(sizeof(int64_t) * (size_t)(ne + 1)), *i = malloc(sizeof(int64_t) * (size_t)(nind > 0 ? nind : 1));
  int64_t *ep = malloc(sizeof(int64_t) * (size_t)(ne > 0 ? ne : 1)), *npv = malloc(sizeof(int64_t) * (size_t)(nn > 0 ? nn : 1));
  int rc = -1;
  if (p && i && ep && npv) {
    for (int e = 0; e <= ne; ++e) p[e] = eptr[e];
    for (int64_t k = 0; k < nind; ++k) i[k] = eind[k];
    rc = METIS_PartMeshNodal(&ne64, &nn64, p, i, NULL, NULL, &np64, NULL, NULL, &objval, ep, npv);
    for (int e = 0; e < ne; ++e) epart[e] = (int)ep[e];
    if (objval_out) *objval_out = objval;
  }
  free(p); free(i); free(ep); free(npv);
  return rc;
}
