"""Synthetic benchmark inputs of arbitrary resolution: the reference's C12 baroclinic fixture (grid metric terms and
analytic Jablonowski-Williamson state of all 6 tiles, tests/golden/c12_6rank_*.npz, produced by the unmodified
reference) resampled bilinearly to C<N> and cut into layout x layout subdomains.

Grid generation and the analytic initial state are out of the hot-path scope (reused from the
reference) and the reference cannot run on the GPU box, so bench.py needs inputs of the right SHAPE that keep the
dycore in its normal regime: every tile keeps its edge / corner metric structure (edge rows map to edge rows, per
staggering), winds / temperature / pressure fields keep their magnitudes, upwind directions and limiter branches
vary in space as in the real case; cell-scale noise (NOISE below) keeps neighbouring cells from taking identical
limiter branches.  Cell sizes stay those of C12, so Courant numbers are ~N/12 times smaller than
the true C<N> case — the arithmetic executed per cell is the same.  Not used by any parity test.
"""
import os

import numpy as np
import torch

GOLDEN = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")
N0 = 12  # resolution of the fixture

# (x interface?, y interface?) of every horizontally staggered input; everything else is cell centred
STAGGER = {
    "grid.dx": (0, 1), "grid.rdx": (0, 1), "grid.dy": (1, 0), "grid.rdy": (1, 0), "grid.dxc": (1, 0), "grid.rdxc": (1, 0),
    "grid.dyc": (0, 1), "grid.rdyc": (0, 1), "grid.rarea_c": (1, 1), "grid.cosa": (1, 1), "grid.rsina": (1, 1),
    "grid.fC": (1, 1), "grid.lon": (1, 1), "grid.lat": (1, 1), "grid.cosa_u": (1, 0), "grid.sina_u": (1, 0),
    "grid.rsin_u": (1, 0), "grid.cosa_v": (0, 1), "grid.sina_v": (0, 1), "grid.rsin_v": (0, 1), "grid.edge_w": (1, 1),
    "grid.edge_e": (1, 1), "grid.edge_s": (1, 1), "grid.edge_n": (1, 1), "damp.del6_u": (0, 1), "damp.del6_v": (1, 0),
    "damp.divg_u": (0, 1), "damp.divg_v": (1, 0),
    "u": (0, 1), "v": (1, 0), "uc": (1, 0), "vc": (0, 1),
}


def _coords(n_tile, interface):
    """source coordinate in the 19-wide fixture arrays for every index of an (n_tile+7)-wide tile array.
    Halo rows and the rows on the tile edge are copied exactly (edge metric terms hold special values that
    must not be blended into their neighbours); the rows in between are stretched linearly."""
    t = np.arange(n_tile + 7, dtype=np.float64)
    last_src = 3 + N0 if interface else 3 + N0 - 1          # source index of the last compute row
    last_tgt = 3 + n_tile if interface else 3 + n_tile - 1
    s = 4.0 + (t - 4.0) * (last_src - 1 - 4.0) / (last_tgt - 1 - 4.0)
    s = np.where(t <= 3, t, s)
    s = np.where(t >= last_tgt, t - (n_tile - N0), s)
    return np.clip(s, 0.0, 18.0)


def resample_tile(a, n_tile, stag, device):
    """a: fixture array (19, 19[, nk]) of one tile -> torch tensor ([nk,] n_tile+7, n_tile+7) laid out [k][j][i]"""
    t = torch.from_numpy(np.ascontiguousarray(a, dtype=np.float64)).to(device)
    if t.ndim == 2:
        t = t.T.unsqueeze(0)            # [1][j][i]
    else:
        t = t.permute(2, 1, 0)          # [k][j][i]
    si = torch.from_numpy(_coords(n_tile, stag[0])).to(device)
    sj = torch.from_numpy(_coords(n_tile, stag[1])).to(device)
    i0 = si.floor().long().clamp(0, 17)
    j0 = sj.floor().long().clamp(0, 17)
    wi = (si - i0).view(1, 1, -1)
    wj = (sj - j0).view(1, -1, 1)
    a00 = t[:, j0][:, :, i0]
    a01 = t[:, j0][:, :, i0 + 1]
    a10 = t[:, j0 + 1][:, :, i0]
    a11 = t[:, j0 + 1][:, :, i0 + 1]
    # exact copies where the weights are 0 (edge rows / halo): junk markers in unused rows must not be blended in
    top = torch.where(wi == 0, a00, a00 * (1 - wi) + a01 * wi)
    bot = torch.where(wi == 0, a10, a10 * (1 - wi) + a11 * wi)
    out = torch.where(wj == 0, top, top * (1 - wj) + bot * wj)
    return out


def _load(kind):
    z = np.load(os.path.join(GOLDEN, "c12_6rank_%s.npz" % kind))
    out = [dict() for _ in range(6)]
    for key in z.files:
        r, name = key.split(":", 1)
        a = z[key]
        if kind != "grid" and a.ndim == 3:
            full = np.zeros((19, 19, a.shape[2]))
            full[3:16, 3:16] = a
            a = full
        out[int(r)][name] = a
    return out


# Small-scale structure on top of the (piecewise-linear) resampled fields: amplitude per field, absolute or relative.
# Without it neighbouring cells see identical slopes, so the PPM limiters and upwind selects of a whole warp take the
# same branch, which flatters instruction-bound kernels.
NOISE = {"u": (0.2, False), "v": (0.2, False), "w": (0.02, False), "pt": (0.05, False), "delp": (1.0e-4, True),
         "qvapor": (1.0e-3, True)}


def cell_noise(shape, tile, field_id, device):
    """Deterministic values in [-1, 1) from the integer position (tile, k, j, i) in the TILE's array — independent of the
    layout and of which GPU holds the cell, and bit-identical on every device (integer arithmetic only)."""
    nk, nj, ni = shape
    k = torch.arange(nk, device=device, dtype=torch.int64).view(-1, 1, 1)
    j = torch.arange(nj, device=device, dtype=torch.int64).view(1, -1, 1)
    i = torch.arange(ni, device=device, dtype=torch.int64).view(1, 1, -1)
    h = (i * 73856093 + j * 19349663 + k * 83492791 + tile * 2654435761 + field_id * 40503) & 0x7FFFFFFF
    for _ in range(2):
        h = (h * 1103515245 + 12345) & 0x7FFFFFFF
        h = h ^ (h >> 13)
    return h.to(torch.float64) / float(1 << 30) - 1.0


def make_engine(n_tile, layout, npz=79, device="cuda:0", comm=None, noise=True, **cfg):
    """A DycoreEngine for C<n_tile> with `layout` x `layout` subdomains per tile, filled with resampled inputs.
    comm: pace_b200.distributed.TorchComm — this process then holds only its block of subdomains.
    noise: add cell-scale perturbations (NOISE) to the prognostic fields."""
    from .distributed import Placement
    from .engine import STATE_2D, STATE_3D, DycoreEngine
    from .hostprep import a2b_corner_coefficients

    grid6, init6 = _load("grid"), _load("init")
    n = n_tile // layout
    mine = Placement(6 * layout * layout, comm.world_size if comm is not None else 1).ranks_of(
        comm.rank if comm is not None else 0)
    slot = {r: s for s, r in enumerate(mine)}
    total = len(mine)
    # DycoreEngine wants per-subdomain grid dicts for the small things (columns, scalars); the big 2-D arrays are
    # written straight into its device tensors below, so hand it light-weight dicts of the right shape
    stub = {}
    for k, v in grid6[0].items():
        if np.ndim(v) == 2:
            stub[k] = np.zeros((n + 7, n + 7))
        elif np.ndim(v) == 1 and k in ("grid.edge_s", "grid.edge_n"):
            stub[k] = np.zeros(n + 7)
        elif np.ndim(v) == 3:
            continue
        else:
            stub[k] = v
    eng = DycoreEngine(n_tile, layout, npz, [stub] * total, comm=comm, device=device, **cfg)
    core = eng.core
    dev = core.device

    def scatter(name, tile_tensor, tile):
        dst = core.t[name]
        for jy in range(layout):
            for ix in range(layout):
                r = slot.get(tile * layout * layout + jy * layout + ix)
                if r is None:
                    continue
                sub = tile_tensor[:, jy * n:jy * n + n + 7, ix * n:ix * n + n + 7]
                if dst.ndim == 4:
                    dst[r, :sub.shape[0], :, :n + 7] = sub
                else:
                    dst[r, :, :n + 7] = sub[0]

    coef = np.zeros((total, 12))
    for tile in range(6):
        if not any(r // (layout * layout) == tile for r in mine):
            continue
        g = grid6[tile]
        for name, a in g.items():
            if np.ndim(a) == 2:
                t = resample_tile(a, n_tile, STAGGER.get(name, (0, 0)), dev)
                scatter(name, t, tile)
            elif name in ("grid.edge_s", "grid.edge_n"):
                a2 = np.repeat(np.asarray(a)[:, None], 19, axis=1)
                scatter(name, resample_tile(a2, n_tile, (1, 1), dev), tile)
        # a2b_ord4 corner-extrapolation weights: those of the fixture tile (ratios of great-circle distances,
        # resolution independent to first order); longitudes cannot be interpolated across the date line
        tile_coef = a2b_corner_coefficients(g["grid.lon"], g["grid.lat"], g["grid.lon_agrid"], g["grid.lat_agrid"], N0, N0)
        for r, sl in slot.items():
            if r // (layout * layout) == tile:
                coef[sl] = tile_coef
        s = init6[tile]
        for name in STATE_3D + STATE_2D:
            a = s[name]
            if not np.any(a):
                continue
            t = resample_tile(a, n_tile, STAGGER.get(name, (0, 0)), dev)
            if noise and name in NOISE and t.shape[0] > 1:
                amp, relative = NOISE[name]
                r = cell_noise(tuple(t.shape), tile, sorted(NOISE).index(name), dev)
                t = t * (1.0 + amp * r) if relative else t + amp * r
            scatter(name, t, tile)
        scatter("ad.zs", resample_tile(np.asarray(s["phis"]) / 9.80665, n_tile, (0, 0), dev), tile)
    core.set_column("grid.a2b_corner_coef", coef.reshape(-1))
    return eng


def setup_coupling(eng, n_tile, layout):
    """Registers the dycore <-> physics coupling fields of `eng` (DycoreEngine.setup_coupling) with DriverGridData arrays
    resampled from the coupling fixture (tests/golden/c12_6rank_coupling.npz), for bench.py --workload coupling_c384."""
    from .engine import DRIVER_GRID

    z = np.load(os.path.join(GOLDEN, "c12_6rank_coupling.npz"))
    n = n_tile // layout
    core = eng.core
    dev = core.device
    eng.setup_coupling([{name: np.zeros((n + 7, n + 7)) for name in DRIVER_GRID}] * len(eng.ranks))
    slot = {r: s for s, r in enumerate(eng.ranks)}
    for tile in range(6):
        if not any(r // (layout * layout) == tile for r in eng.ranks):
            continue
        for name in DRIVER_GRID:
            a = np.nan_to_num(z["r%d:grid:%s" % (tile, name)])
            if a.ndim == 1:
                a = np.repeat(a[:, None], 19, axis=1)
            t = resample_tile(a, n_tile, (0, 0), dev)
            dst = core.t["dgrid." + name]
            for jy in range(layout):
                for ix in range(layout):
                    r = slot.get(tile * layout * layout + jy * layout + ix)
                    if r is not None:
                        dst[r, :, :n + 7] = t[0, jy * n:jy * n + n + 7, ix * n:ix * n + n + 7]
