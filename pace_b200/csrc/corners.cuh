// Cube-corner halo fills (reference NDSL/ndsl/stencils/corners.py).
//
// At the four corners of a cube tile the 3x3 halo block has no neighbour tile; before a 1-D (x or y)
// difference operator crosses the corner these cells are filled from the adjacent halo strip, rotated.
// All of these touch at most 3 cells per corner per level, so each family is one small launch that
// executes a list of independent "ops" (one CUDA block per (subdomain, level), one thread per cell).
#pragma once
#include "core.hpp"

namespace pace {

struct CornerOp {
  double* q;          // field written
  const double* qc;   // field read (may alias q: the cells read are never cells written)
  int ncells;         // 2 or 3 cells per corner
  int ydir;           // 0: fill_corners_*_x, 1: fill_corners_*_y
  double sw, se, nw, ne;
};
struct CornerOps {
  CornerOp op[8];
  int n;
  int nk;
};

// fill_corners_{2,3}cells_mult_{x,y} (corners.py:129-304), in absolute indices:
//   x:  q(is-n, js-1) = sw*qc(is-1, js-1+n)   q(ie+n, js-1) = se*qc(ie+1, js-1+n)
//       q(is-n, je+1) = nw*qc(is-1, je+1-n)   q(ie+n, je+1) = ne*qc(ie+1, je+1-n)
//   y:  q(is-1, js-n) = sw*qc(is-1+n, js-1)   q(ie+1, js-n) = se*qc(ie+1-n, js-1)
//       q(is-1, je+n) = nw*qc(is-1+n, je+1)   q(ie+1, je+n) = ne*qc(ie+1-n, je+1)
__device__ __forceinline__ void corner_fill_cell(const Geo& g, const CornerOp& o, int t, int k, int c, int n) {
  const unsigned e = edge_of(g, t);
  bool W = e & EDGE_W, E = e & EDGE_E, S = e & EDGE_S, N = e & EDGE_N;
  int is = g.is, ie = g.ie, js = g.js, je = g.je;
  int wi, wj, ri, rj;
  double m;
  bool on;
  if (!o.ydir) {
    switch (c) {
      case 0: on = W && S; wi = is - n; wj = js - 1; ri = is - 1; rj = js - 1 + n; m = o.sw; break;
      case 1: on = E && S; wi = ie + n; wj = js - 1; ri = ie + 1; rj = js - 1 + n; m = o.se; break;
      case 2: on = W && N; wi = is - n; wj = je + 1; ri = is - 1; rj = je + 1 - n; m = o.nw; break;
      default: on = E && N; wi = ie + n; wj = je + 1; ri = ie + 1; rj = je + 1 - n; m = o.ne; break;
    }
  } else {
    switch (c) {
      case 0: on = W && S; wi = is - 1; wj = js - n; ri = is - 1 + n; rj = js - 1; m = o.sw; break;
      case 1: on = E && S; wi = ie + 1; wj = js - n; ri = ie + 1 - n; rj = js - 1; m = o.se; break;
      case 2: on = W && N; wi = is - 1; wj = je + n; ri = is - 1 + n; rj = je + 1; m = o.nw; break;
      default: on = E && N; wi = ie + 1; wj = je + n; ri = ie + 1 - n; rj = je + 1; m = o.ne; break;
    }
  }
  if (!on) return;
  long long base = (long long)t * g.st3 + (long long)k * g.sk;
  o.q[base + wj * g.sj + wi] = m * o.qc[base + rj * g.sj + ri];
}

#ifndef PACE_HOST_EMU
static __global__ void k_corner_ops(Geo g, CornerOps ops) {
  int k = blockIdx.x % ops.nk, t = blockIdx.x / ops.nk;
  int id = threadIdx.x;            // op*12 + corner*3 + (n-1)
  int io = id / 12, c = (id % 12) / 3, n = id % 3 + 1;
  if (io >= ops.n || n > ops.op[io].ncells) return;
  corner_fill_cell(g, ops.op[io], t, k, c, n);
}
inline void run_corner_ops(cudaStream_t s, const Geo& g, const CornerOps& ops) {
  ++launch_counter();
  k_corner_ops<<<ops.nk * g.nt, 96, 0, s>>>(g, ops);
}
#else
inline void run_corner_ops(cudaStream_t, const Geo& g, const CornerOps& ops) {
  for (int t = 0; t < g.nt; ++t)
    for (int k = 0; k < ops.nk; ++k)
      for (int io = 0; io < ops.n; ++io)
        for (int c = 0; c < 4; ++c)
          for (int n = 1; n <= ops.op[io].ncells; ++n) corner_fill_cell(g, ops.op[io], t, k, c, n);
}
#endif

}  // namespace pace

namespace pace {
// Virtual corner fill for reads: the cell that fill_corners_2cells_x / _y (mult 1) would have copied
// into (i,j).  Lets a consumer read q "as if filled" without materialising the filled copy.
struct EdgeInfo {
  int is, ie, js, je;
  bool W, E, S, N;
};
__device__ __forceinline__ EdgeInfo edge_info(const Geo& g, const Pt& p) {
  return EdgeInfo{g.is, g.ie, g.js, g.je, p.W, p.E, p.S, p.N};
}
__device__ __forceinline__ void xfill_src(const EdgeInfo& e, int nc, int& i, int& j) {
  if (e.S && j == e.js - 1) {
    if (e.W && i < e.is && i >= e.is - nc) { int n = e.is - i; i = e.is - 1; j = e.js - 1 + n; }
    else if (e.E && i > e.ie && i <= e.ie + nc) { int n = i - e.ie; i = e.ie + 1; j = e.js - 1 + n; }
  } else if (e.N && j == e.je + 1) {
    if (e.W && i < e.is && i >= e.is - nc) { int n = e.is - i; i = e.is - 1; j = e.je + 1 - n; }
    else if (e.E && i > e.ie && i <= e.ie + nc) { int n = i - e.ie; i = e.ie + 1; j = e.je + 1 - n; }
  }
}
__device__ __forceinline__ void yfill_src(const EdgeInfo& e, int nc, int& i, int& j) {
  if (e.W && i == e.is - 1) {
    if (e.S && j < e.js && j >= e.js - nc) { int n = e.js - j; j = e.js - 1; i = e.is - 1 + n; }
    else if (e.N && j > e.je && j <= e.je + nc) { int n = j - e.je; j = e.je + 1; i = e.is - 1 + n; }
  } else if (e.E && i == e.ie + 1) {
    if (e.S && j < e.js && j >= e.js - nc) { int n = e.js - j; j = e.js - 1; i = e.ie + 1 - n; }
    else if (e.N && j > e.je && j <= e.je + nc) { int n = j - e.je; j = e.je + 1; i = e.ie + 1 - n; }
  }
}
}  // namespace pace

namespace pace {
// copy_corners_x / copy_corners_y (corners.py:307-424; same tables in delnflux.py copy_corners_*_nord):
// each 3x3 corner block of the halo is overwritten with the rotated neighbouring halo strip.  In
// absolute indices, with a = distance from the tile edge in x (1..3), b = distance in y (1..3):
//   x-copy:  (is-a, js-b) <- (is-b, js-1+a)    (ie+a, js-b) <- (ie+b, js-1+a)
//            (is-a, je+b) <- (is-b, je+1-a)    (ie+a, je+b) <- (ie+b, je+1-a)
//   y-copy:  (is-a, js-b) <- (is-1+b, js-a)    (ie+a, js-b) <- (ie+1-b, js-a)
//            (is-a, je+b) <- (is-1+b, je+a)    (ie+a, je+b) <- (ie+1-b, je+a)
// Sources are never corner cells, so a consumer may equally read "as if copied" (virtual copy).
__device__ __forceinline__ void copyx_src(const EdgeInfo& e, int& i, int& j) {
  if (e.S && j < e.js) {
    if (e.W && i < e.is) { int a = e.is - i, b = e.js - j; i = e.is - b; j = e.js - 1 + a; }
    else if (e.E && i > e.ie) { int a = i - e.ie, b = e.js - j; i = e.ie + b; j = e.js - 1 + a; }
  } else if (e.N && j > e.je) {
    if (e.W && i < e.is) { int a = e.is - i, b = j - e.je; i = e.is - b; j = e.je + 1 - a; }
    else if (e.E && i > e.ie) { int a = i - e.ie, b = j - e.je; i = e.ie + b; j = e.je + 1 - a; }
  }
}
__device__ __forceinline__ void copyy_src(const EdgeInfo& e, int& i, int& j) {
  if (e.S && j < e.js) {
    if (e.W && i < e.is) { int a = e.is - i, b = e.js - j; i = e.is - 1 + b; j = e.js - a; }
    else if (e.E && i > e.ie) { int a = i - e.ie, b = e.js - j; i = e.ie + 1 - b; j = e.js - a; }
  } else if (e.N && j > e.je) {
    if (e.W && i < e.is) { int a = e.is - i, b = j - e.je; i = e.is - 1 + b; j = e.je + a; }
    else if (e.E && i > e.ie) { int a = i - e.ie, b = j - e.je; i = e.ie + 1 - b; j = e.je + a; }
  }
}
// value of f at (p.i+di, p.j+dj, p.k) as seen after copy_corners_x / _y
__device__ __forceinline__ double rd_cx(const double* f, const Pt& p, const EdgeInfo& e, int di, int dj) {
  int i = p.i + di, j = p.j + dj;
  copyx_src(e, i, j);
  return f[p.o + (i - p.i) + (j - p.j) * p.sj];
}
__device__ __forceinline__ double rd_cy(const double* f, const Pt& p, const EdgeInfo& e, int di, int dj) {
  int i = p.i + di, j = p.j + dj;
  copyy_src(e, i, j);
  return f[p.o + (i - p.i) + (j - p.j) * p.sj];
}

// materialised copy_corners_x / _y of `q` in place, levels [0,nk)
void copy_corners(Core& c, double* q, bool ydir, int nk);
}  // namespace pace
