// plan_hex.cpp -- integer pass for the fused hex8 kernel (assembly variant 2).
//
// Proves that a hex8 mesh is a subset of a logical (i,j,k) grid with consistently oriented
// elements (node order of ShapeFuncdHex, src/m_elems.F90:574-579) by a breadth-first walk
// over shared nodes -- node NUMBERING is irrelevant, only connectivity is used -- and emits
//   * dense logical maps  node_at[k][j][i], elem_at[k][j][i]  (-1 where absent),
//   * per node the position k of each of its 27 canonical neighbours (dx,dy,dz) in that
//     node's sorted block row (255 where the neighbour is not coupled).
// With those the kernel needs no per-element slot map: a node's row is the sum over its
// <= 8 corner elements in canonical octant order.  Meshes that fail any check keep using
// the two-pass path (variant 1); nothing is approximated.
#include <algorithm>
#include <queue>

#include "femx_internal.h"

namespace femx {

static const int kSx[8] = {0, 1, 1, 0, 0, 1, 1, 0};
static const int kSy[8] = {0, 0, 1, 1, 0, 0, 1, 1};
static const int kSz[8] = {0, 0, 0, 0, 1, 1, 1, 1};

int build_hex_plan(const Plan &P, const int *etype, const int *eptr, HexPlan *H, std::string *why) {
  H->ok = false;
  const int nels = P.nels, nn = P.nlnds;
  if (P.pdim != 3 || nels == 0) { *why = "not a 3-D mesh"; return 0; }
  for (int e = 0; e < nels; ++e)
    if (etype[e] != FEMX_HEX) { *why = "mesh holds non-hex8 elements"; return 0; }
  // breadth-first assignment of logical coordinates
  const int UNSET = INT32_MIN;
  std::vector<int32_t> ci(nn, UNSET), cj(nn, UNSET), ck(nn, UNSET);
  std::vector<int32_t> ei(nels, UNSET), ej(nels), ek(nels);
  int next_offset = 0;
  for (int seed = 0; seed < nels; ++seed) {
    if (ei[seed] != UNSET) continue;
    std::queue<int> q;
    ei[seed] = next_offset; ej[seed] = 0; ek[seed] = 0;
    q.push(seed);
    int comp_max_i = next_offset;
    while (!q.empty()) {
      const int e = q.front();
      q.pop();
      const int32_t *c = &P.conn0[eptr[e]];
      for (int b = 0; b < 8; ++b) {
        const int r = c[b];
        const int x = ei[e] + kSx[b], y = ej[e] + kSy[b], z = ek[e] + kSz[b];
        if (ci[r] == UNSET) {
          ci[r] = x; cj[r] = y; ck[r] = z;
          comp_max_i = std::max(comp_max_i, x);
          for (int64_t p = P.n2e_ptr[r]; p < P.n2e_ptr[r + 1]; ++p) {
            const int e2 = P.n2e_elem[p], a2 = P.n2e_a[p];
            if (ei[e2] == UNSET) {
              ei[e2] = x - kSx[a2]; ej[e2] = y - kSy[a2]; ek[e2] = z - kSz[a2];
              q.push(e2);
            }
          }
        } else if (ci[r] != x || cj[r] != y || ck[r] != z) {
          *why = "elements are not consistently oriented on a logical grid";
          return 0;
        }
      }
    }
    next_offset = comp_max_i + 3;  // keep disconnected components apart
  }
  int mn[3] = {INT32_MAX, INT32_MAX, INT32_MAX}, mx[3] = {INT32_MIN, INT32_MIN, INT32_MIN};
  for (int r = 0; r < nn; ++r) {
    if (ci[r] == UNSET) { *why = "a node belongs to no element"; return 0; }
    mn[0] = std::min(mn[0], ci[r]); mn[1] = std::min(mn[1], cj[r]); mn[2] = std::min(mn[2], ck[r]);
    mx[0] = std::max(mx[0], ci[r]); mx[1] = std::max(mx[1], cj[r]); mx[2] = std::max(mx[2], ck[r]);
  }
  const int64_t NI = mx[0] - mn[0] + 1, NJ = mx[1] - mn[1] + 1, NK = mx[2] - mn[2] + 1;
  if (NI * NJ * NK > 4 * (int64_t)nn + 4096) { *why = "logical bounding grid is too sparse"; return 0; }
  H->NI = (int)NI; H->NJ = (int)NJ; H->NK = (int)NK;
  H->node_at.assign(NI * NJ * NK, -1);
  H->elem_at.assign(NI * NJ * NK, -1);  // element (i,j,k) spans nodes i..i+1; same extents for simplicity
  auto at = [&](int i, int j, int k) { return ((int64_t)k * NJ + j) * NI + i; };
  for (int r = 0; r < nn; ++r) {
    int32_t &slot = H->node_at[at(ci[r] - mn[0], cj[r] - mn[1], ck[r] - mn[2])];
    if (slot != -1) { *why = "two nodes share one logical position"; return 0; }
    slot = r;
  }
  for (int e = 0; e < nels; ++e) {
    int32_t &slot = H->elem_at[at(ei[e] - mn[0], ej[e] - mn[1], ek[e] - mn[2])];
    if (slot != -1) { *why = "two elements share one logical cell"; return 0; }
    slot = e;
  }
  // canonical neighbour -> block position; every block of the row must be reached this way
  H->kmap.assign((size_t)nn * 32, 255);
  for (int r = 0; r < nn; ++r) {
    const int i = ci[r] - mn[0], j = cj[r] - mn[1], k = ck[r] - mn[2];
    const int32_t *b0 = &P.bcol[P.bptr[r]], *b1 = &P.bcol[P.bptr[r + 1]];
    int found = 0;
    for (int dz = -1; dz <= 1; ++dz)
      for (int dy = -1; dy <= 1; ++dy)
        for (int dx = -1; dx <= 1; ++dx) {
          const int x = i + dx, y = j + dy, z = k + dz;
          if (x < 0 || y < 0 || z < 0 || x >= NI || y >= NJ || z >= NK) continue;
          const int c = H->node_at[at(x, y, z)];
          if (c < 0) continue;
          const int32_t *it = std::lower_bound(b0, b1, c);
          if (it == b1 || *it != c) continue;  // present on the grid but not coupled through an element
          H->kmap[(size_t)r * 32 + (dz + 1) * 9 + (dy + 1) * 3 + (dx + 1)] = (uint8_t)(it - b0);
          ++found;
        }
    if (found != (int)(b1 - b0) || b1 - b0 > 27) { *why = "a node couples to nodes outside its 27-point stencil"; return 0; }
  }
  H->ok = true;
  return 1;
}

}  // namespace femx
