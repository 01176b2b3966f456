"""Cubed-sphere halo-exchange plan: which cell of which subdomain fills each halo cell, as affine index maps.

Replaces the reference's CubedSpherePartitioner / Boundary / HaloUpdater index logic
(NDSL/ndsl/comm/partitioner.py:373-614, comm/_boundary_utils.py:57-95, halo/rotate.py:4-50,
halo/updater.py:217-303, 385-515) with closed-form maps — no per-element index arrays:

  A point of tile t is addressed by doubled coordinates (X, Y) = (2x, 2y): cell centres are odd, interfaces
  even, so one map serves every staggering.  Leaving tile t through an edge lands on the neighbour tile with

      even t:  E -> t+1: (X-2N, Y)      S -> t-1: (X, Y+2N)     W -> t-2: (2N-Y, 2N+X)   N -> t+2: (Y-2N, 2N-X)
      odd  t:  W -> t-1: (X+2N, Y)      N -> t+1: (X, Y-2N)     E -> t+2: (2N-Y, X-2N)   S -> t-2: (2N+Y, 2N-X)

  (N = cells per tile edge).  A rotated edge swaps the roles of the x- and y-staggered components of a vector
  pair and flips one sign: with L the linear part of the map, (u, v)_here = L^T (u', v')_there.

A halo update fills, for every subdomain, 4 edge strips and 4 corner blocks; each of these 8 regions has exactly
one sending subdomain (none at the 8 cube corners).  Every region becomes one `HaloOp`.

The plan is pure index logic; it is checked bit-exactly against source maps produced by the reference itself
(tests/golden/halo_maps.npz).  The device side (csrc/halo.cu) executes the ops of all subdomains resident on a
GPU in one launch per field group; ops whose source lives on another GPU go through pack / NCCL / unpack.
"""
from dataclasses import dataclass
from typing import List, Optional, Tuple

import numpy as np

H = 3  # allocated halo width of every field (NDSL N_HALO_DEFAULT)


@dataclass
class HaloOp:
    dst_rank: int
    src_rank: int
    comp: int              # 0: scalar or x-component array is written, 1: y-component array is written
    src_comp: int          # which component array is read
    sign: float
    i0: int                # destination box in array indices (i0..i0+ni, j0..j0+nj)
    j0: int
    ni: int
    nj: int
    # source array index = (aii*i + aij*j + bi, aji*i + ajj*j + bj) for destination array index (i, j)
    aii: int
    aij: int
    bi: int
    aji: int
    ajj: int
    bj: int


def _edge_map(tile: int, X: int, Y: int, n2: int) -> Optional[Tuple[int, Tuple[int, int, int, int, int, int]]]:
    """Map doubled tile coordinates that left `tile` through exactly one edge.
    Returns (neighbour tile, (axx, axy, bx, ayx, ayy, by)) with X' = axx*X + axy*Y + bx, ... or None inside."""
    west, east, south, north = X < 0, X > n2, Y < 0, Y > n2
    if not (west or east or south or north):
        return None
    if tile % 2 == 0:
        if east:
            return (tile + 1) % 6, (1, 0, -n2, 0, 1, 0)
        if south:
            return (tile - 1) % 6, (1, 0, 0, 0, 1, n2)
        if west:
            return (tile - 2) % 6, (0, -1, n2, 1, 0, n2)
        return (tile + 2) % 6, (0, 1, -n2, -1, 0, n2)
    if west:
        return (tile - 1) % 6, (1, 0, n2, 0, 1, 0)
    if north:
        return (tile + 1) % 6, (1, 0, 0, 0, 1, -n2)
    if east:
        return (tile + 2) % 6, (0, -1, n2, 1, 0, -n2)
    return (tile - 2) % 6, (0, 1, n2, -1, 0, n2)


class HaloPlanner:
    """Plans for a cube of 6 tiles x (layout x layout) square subdomains of n x n cells."""

    def __init__(self, n_tile: int, layout: int):
        if n_tile % layout:
            raise ValueError("tile size must be divisible by the layout")
        self.N = n_tile
        self.L = layout
        self.n = n_tile // layout
        self.total_ranks = 6 * layout * layout

    def rank_of(self, tile, ix, jy):
        return tile * self.L * self.L + jy * self.L + ix

    def position(self, rank):
        tile, r = divmod(rank, self.L * self.L)
        jy, ix = divmod(r, self.L)
        return tile, ix, jy

    def edge_flags(self, rank):
        """bit0 W, bit1 E, bit2 S, bit3 N: the subdomain touches that edge of its tile."""
        _, ix, jy = self.position(rank)
        return (ix == 0) | ((ix == self.L - 1) << 1) | ((jy == 0) << 2) | ((jy == self.L - 1) << 3)

    def _region_ops(self, rank, xi: bool, yi: bool, n_points: int, comp: int, vector: bool) -> List[HaloOp]:
        """Ops filling the halo of one array of `rank` whose x / y axes are interface (xi / yi) or centred."""
        n, n2 = self.n, 2 * self.N
        tile, ix, jy = self.position(rank)
        ex, ey = n + int(xi), n + int(yi)          # compute extent in array points
        ox, oy = ix * n, jy * n                    # subdomain origin in tile cells
        sx, sy = (0 if xi else 1), (0 if yi else 1)  # doubled-coordinate stagger offsets
        ops = []
        strips_x = [(-n_points, n_points, -1), (0, ex, 0), (ex, n_points, 1)]  # (start, size, side) in local idx
        strips_y = [(-n_points, n_points, -1), (0, ey, 0), (ey, n_points, 1)]
        for (li0, lni, sidex) in strips_x:
            for (lj0, lnj, sidey) in strips_y:
                if sidex == 0 and sidey == 0:
                    continue
                # representative cell centre of the region decides the sending subdomain
                ci = -1 if sidex < 0 else (n if sidex > 0 else 0)
                cj = -1 if sidey < 0 else (n if sidey > 0 else 0)
                Xc, Yc = 2 * (ox + ci) + 1, 2 * (oy + cj) + 1
                out_x, out_y = (Xc < 0 or Xc > n2), (Yc < 0 or Yc > n2)
                if out_x and out_y:
                    continue  # cube corner: nobody owns it
                m = _edge_map(tile, Xc, Yc, n2)
                if m is None:
                    stile, (axx, axy, bx, ayx, ayy, by) = tile, (1, 0, 0, 0, 1, 0)
                else:
                    stile, (axx, axy, bx, ayx, ayy, by) = m
                Xs, Ys = axx * Xc + axy * Yc + bx, ayx * Xc + ayy * Yc + by
                six, sjy = (Xs // 2) // n, (Ys // 2) // n
                src_rank = self.rank_of(stile, six, sjy)
                sox, soy = six * n, sjy * n
                rotated = axx == 0
                # staggering of the array that is read: rotation swaps x / y staggering
                src_comp, sign = comp, 1.0
                if rotated:
                    s_sx, s_sy = sy, sx
                    if vector:
                        src_comp = 1 - comp
                        # (u, v)_here = L^T (u', v')_there with L = [[axx, axy], [ayx, ayy]]
                        sign = float(ayx) if comp == 0 else float(axy)
                else:
                    s_sx, s_sy = sx, sy
                # destination array index (i, j) -> doubled tile coords X = 2*(ox + i - H) + sx
                # source array index i' = (X' - s_sx)/2 - sox + H
                def lin(a_x, a_y, b, s_st, s_o):
                    # X' = a_x*X + a_y*Y + b ;  i' = (X' - s_st)/2 - s_o + H, affine in (i, j)
                    c0 = a_x * (2 * (ox - H) + sx) + a_y * (2 * (oy - H) + sy) + b - s_st
                    assert c0 % 2 == 0
                    return a_x, a_y, c0 // 2 - s_o + H

                aii, aij, bi = lin(axx, axy, bx, s_sx, sox)
                aji, ajj, bj = lin(ayx, ayy, by, s_sy, soy)
                ops.append(HaloOp(rank, src_rank, comp, src_comp, sign, li0 + H, lj0 + H, lni, lnj,
                                  aii, aij, bi, aji, ajj, bj))
        return ops

    # -- public plans ------------------------------------------------------------------------------
    def scalar_plan(self, xi=False, yi=False, n_points=3, ranks=None) -> List[HaloOp]:
        ops = []
        for r in (range(self.total_ranks) if ranks is None else ranks):
            ops += self._region_ops(r, xi, yi, n_points, 0, False)
        return ops

    def vector_plan(self, grid: str, n_points=3, ranks=None) -> List[HaloOp]:
        """grid 'd': x-component on [x, y_interface], y-component on [x_interface, y] (u, v);
        grid 'c': x-component on [x_interface, y], y-component on [x, y_interface] (uc, vc)."""
        stag = {"d": ((False, True), (True, False)), "c": ((True, False), (False, True))}[grid]
        ops = []
        for r in (range(self.total_ranks) if ranks is None else ranks):
            ops += self._region_ops(r, stag[0][0], stag[0][1], n_points, 0, True)
            ops += self._region_ops(r, stag[1][0], stag[1][1], n_points, 1, True)
        return ops

    def interface_sync_plan(self, ranks=None) -> List[HaloOp]:
        """synchronize_vector_interfaces(u, v) on the D grid (NDSL halo/updater.py:385-515): the north row of
        u and the east column of v are overwritten by the neighbour's south row / west column."""
        n, n2 = self.n, 2 * self.N
        ops = []
        for rank in (range(self.total_ranks) if ranks is None else ranks):
            tile, ix, jy = self.position(rank)
            ox, oy = ix * n, jy * n
            for comp in (0, 1):
                # comp 0: u[x, yi] row j = n ; comp 1: v[xi, y] column i = n
                sx, sy = (1, 0) if comp == 0 else (0, 1)
                if comp == 0:
                    li0, lj0, lni, lnj = 0, n, n, 1
                    Xc, Yc = 2 * ox + 1, 2 * (oy + n) + 1      # cell just north of the row
                else:
                    li0, lj0, lni, lnj = n, 0, 1, n
                    Xc, Yc = 2 * (ox + n) + 1, 2 * oy + 1      # cell just east of the column
                m = _edge_map(tile, Xc, Yc, n2)
                if m is None:
                    stile, (axx, axy, bx, ayx, ayy, by) = tile, (1, 0, 0, 0, 1, 0)
                else:
                    stile, (axx, axy, bx, ayx, ayy, by) = m
                Xs, Ys = axx * Xc + axy * Yc + bx, ayx * Xc + ayy * Yc + by
                six, sjy = (Xs // 2) // n, (Ys // 2) // n
                src_rank = self.rank_of(stile, six, sjy)
                sox, soy = six * n, sjy * n
                rotated = axx == 0
                src_comp, sign = comp, 1.0
                s_sx, s_sy = sx, sy
                if rotated:
                    s_sx, s_sy = sy, sx
                    src_comp = 1 - comp
                    sign = float(ayx) if comp == 0 else float(axy)

                def lin(a_x, a_y, b, s_st, s_o):
                    c0 = a_x * (2 * (ox - H) + sx) + a_y * (2 * (oy - H) + sy) + b - s_st
                    assert c0 % 2 == 0
                    return a_x, a_y, c0 // 2 - s_o + H

                aii, aij, bi = lin(axx, axy, bx, s_sx, sox)
                aji, ajj, bj = lin(ayx, ayy, by, s_sy, soy)
                ops.append(HaloOp(rank, src_rank, comp, src_comp, sign, li0 + H, lj0 + H, lni, lnj,
                                  aii, aij, bi, aji, ajj, bj))
        return ops


def apply_plan_numpy(ops: List[HaloOp], arrays_x, arrays_y=None):
    """Reference executor on host arrays [rank][i, j(, k)] (tests + CPU multi-process logic checks).
    All reads see the pre-update state (sources are compute-domain cells, never halo cells)."""
    src = [[a.copy() for a in arrays_x], None if arrays_y is None else [a.copy() for a in arrays_y]]
    dst = [arrays_x, arrays_y]
    for op in ops:
        ii, jj = np.meshgrid(np.arange(op.i0, op.i0 + op.ni), np.arange(op.j0, op.j0 + op.nj), indexing="ij")
        si = op.aii * ii + op.aij * jj + op.bi
        sj = op.aji * ii + op.ajj * jj + op.bj
        vals = src[op.src_comp][op.src_rank][si, sj]
        dst[op.comp][op.dst_rank][op.i0:op.i0 + op.ni, op.j0:op.j0 + op.nj] = op.sign * vals if op.sign != 1.0 else vals
