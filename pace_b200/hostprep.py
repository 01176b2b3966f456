"""Host-side, init-time preparation of the small inputs the kernels read: the per-level "column
namelist" of d_sw, the delnflux damping columns and the grid-only constants of a2b_ord4.

Everything here mirrors init-time Python of the reference (cited per function) and uses `math.*`
(the C libm) for transcendentals so results do not depend on NumPy's SIMD dispatch.
"""
import math

import numpy as np


def column_namelist(cfg, npz):
    """pyFV3/stencils/d_sw.py:647-691 get_column_namelist (+ set_low_kvals / vorticity_damping_option)."""
    col = {}
    names = ["ke_bg", "d_con", "nord", "nord_v", "nord_w", "nord_t", "damp_vt", "damp_w", "damp_t", "d2_divg"]
    for n in names:
        col[n] = np.zeros(npz, dtype=np.float64)
    for n in ("ke_bg", "d_con", "nord"):
        col[n][:] = cfg[n]
    col["d2_divg"][:] = min(0.2, cfg["d2_bg"])
    col["nord_v"][:] = min(2, col["nord"][0])
    col["nord_w"][:] = col["nord_v"][0]
    col["nord_t"][:] = col["nord_v"][0]
    col["damp_vt"][:] = cfg["vtdm4"] if cfg["do_vort_damp"] else 0
    col["damp_w"][:] = col["damp_vt"][0]
    col["damp_t"][:] = col["damp_vt"][0]

    def set_low_kvals(k):
        for n in ("nord", "nord_w", "d_con"):
            col[n][k] = 0
        col["damp_w"][k] = col["d2_divg"][k]

    def lowest_kvals(k):
        set_low_kvals(k)
        if cfg["do_vort_damp"]:
            col["nord_v"][k] = 0
            col["damp_vt"][k] = 0.5 * col["d2_divg"][k]

    if npz == 1 or cfg["n_sponge"] < 0:
        col["d2_divg"][0] = cfg["d2_bg"]
    else:
        col["d2_divg"][0] = max(0.01, cfg["d2_bg"], cfg["d2_bg_k1"])
        lowest_kvals(0)
        if cfg["d2_bg_k2"] > 0.01:
            col["d2_divg"][1] = max(cfg["d2_bg"], cfg["d2_bg_k2"])
            lowest_kvals(1)
        if cfg["d2_bg_k2"] > 0.05:
            col["d2_divg"][2] = max(cfg["d2_bg"], 0.2 * cfg["d2_bg_k2"])
            set_low_kvals(2)
    return col


def calc_damp(damp_c, da_min, nord):
    """pyFV3/stencils/delnflux.py:22-38: (damp_c * da_min) ** (nord + 1), per level."""
    return np.array([math.pow(float(d) * float(da_min), float(n) + 1.0) for d, n in zip(damp_c, nord)] + [0.0])


def all_columns(cfg, npz, da_min, da_min_c):
    """Every K-column the library expects under "col.*" (d_sw.py:855-990, fvtp2d.py:168-180)."""
    col = column_namelist(cfg, npz)
    # DelnFluxNoSG rounds its nord column in place (delnflux.py:1069)
    for n in ("nord_v", "nord_w", "nord_t"):
        col[n] = np.round(col[n])
    # [Z_DIM] quantities are allocated with npz+1 entries (NDSL sizer.py:141-154); stencils that run over
    # npz+1 levels (updatedzd's delnflux) read the trailing 0
    out = {"col." + k: np.concatenate([v, [0.0]]) for k, v in col.items()}
    out["col.delnflux_damp_vt"] = calc_damp(col["damp_vt"], da_min_c, col["nord_v"])  # d_sw.py:980-984
    out["col.delnflux_damp_w"] = calc_damp(col["damp_w"], da_min_c, col["nord_w"])    # d_sw.py:985-989
    out["col.delnflux_damp_vt_fv"] = calc_damp(col["damp_vt"], da_min, col["nord_v"])  # DelnFlux in fvtp2d_dp / _tm
    out["col.delnflux_damp_t_fv"] = calc_damp(col["damp_t"], da_min, col["nord_t"])    # DelnFlux in fvtp2d_dp_t
    return out


def cubic_spline_constants(dp_ref, npz):
    """pyFV3/stencils/updatedzd.py:130-167 cubic_spline_interpolation_constants."""
    dp0 = np.asarray(dp_ref, dtype=np.float64)[:npz]
    gk = np.zeros(npz + 1)
    beta = np.zeros(npz + 1)
    gamma = np.zeros(npz + 1)
    gk[0] = dp0[1] / dp0[0]
    beta[0] = gk[0] * (gk[0] + 0.5)
    gamma[0] = (1.0 + gk[0] * (gk[0] + 1.5)) / beta[0]
    gk[1:npz] = dp0[:-1] / dp0[1:]
    for i in range(1, npz):
        beta[i] = 2.0 + 2.0 * gk[i] - gamma[i - 1]
        gamma[i] = gk[i] / beta[i]
    return {"col.dzd_gk": gk, "col.dzd_beta": beta, "col.dzd_gamma": gamma}


def _great_circle_dist(p1a, p1b, p2a, p2b):
    """pyFV3/stencils/a2b_ord4.py:36-40 (x**2.0 evaluates as x*x in the numpy backend)."""
    s = math.sin((p1b - p2b) / 2.0)
    tb = s * s
    s = math.sin((p1a - p2a) / 2.0)
    ta = s * s
    return math.asin(math.sqrt(tb + math.cos(p1b) * math.cos(p2b) * ta)) * 2.0


def a2b_corner_coefficients(lon, lat, lon_agrid, lat_agrid, nx, ny):
    """x1/(x2-x1) of extrap_corner (a2b_ord4.py:43-59) for the 4 tile corners x 3 extrapolations, in the
    library's order [SW, SE, NE, NW][ec1, ec2, ec3] (stencils _sw/_nw/_ne/_se_corner, a2b_ord4.py:62-247)."""
    is_, js, ie1, je1 = 3, 3, 3 + nx, 3 + ny
    # (corner point, [(qa offset, qb offset) x3])
    spec = [
        ((is_, js), [((0, 0), (1, 1)), ((-1, 0), (-2, 1)), ((0, -1), (1, -2))]),
        ((ie1, js), [((-1, 0), (-2, 1)), ((-1, -1), (-2, -2)), ((0, 0), (1, 1))]),
        ((ie1, je1), [((-1, -1), (-2, -2)), ((0, -1), (1, -2)), ((-1, 0), (-2, 1))]),
        ((is_, je1), [((0, -1), (1, -2)), ((-1, -1), (-2, -2)), ((0, 0), (1, 1))]),
    ]
    out = np.zeros(12, dtype=np.float64)
    n = 0
    for (ci, cj), pairs in spec:
        p0a, p0b = float(lon[ci, cj]), float(lat[ci, cj])
        for (a, b) in pairs:
            x1 = _great_circle_dist(float(lon_agrid[ci + a[0], cj + a[1]]), float(lat_agrid[ci + a[0], cj + a[1]]), p0a, p0b)
            x2 = _great_circle_dist(float(lon_agrid[ci + b[0], cj + b[1]]), float(lat_agrid[ci + b[0], cj + b[1]]), p0a, p0b)
            out[n] = x1 / (x2 - x1) if x2 != x1 else 0.0
            n += 1
    return out


def derived_grid_fields(grids, nx, ny):
    """grids: list (one per subdomain) of dicts name -> reference-layout numpy array ("grid.*", "damp.*").
    Returns extra per-subdomain lists / flat arrays to register with the library."""
    extra = {}
    # 1-D I-fields (edge_s, edge_n: a2b_ord4.py qout_y_edge) broadcast to (i, j) so every subdomain has its own
    for name in ("grid.edge_s", "grid.edge_n"):
        extra[name] = [np.repeat(np.asarray(g[name])[:, None], ny + 7, axis=1) for g in grids]
    extra["grid.a2b_corner_coef"] = [np.concatenate([
        a2b_corner_coefficients(g["grid.lon"], g["grid.lat"], g["grid.lon_agrid"], g["grid.lat_agrid"], nx, ny)
        for g in grids
    ])]
    return extra
