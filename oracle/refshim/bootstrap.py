"""TEST INFRASTRUCTURE ONLY — bring the *unmodified* reference (pyFV3 + NDSL + vendored GT4Py numpy
backend under /root/reference) up inside this container so it can be used as the parity oracle and to
generate the golden fixtures in tests/golden/.  Nothing in the product (pace_b200/) imports this.

The reference cannot travel to the GPU box, so this module is only ever used here (fixture generation,
oracle-vs-restatement checks); tests marked `gpu` use the committed fixtures instead.

Recipe follows SURVEY.md Appendix A.  No reference source is copied: missing third-party packages are
replaced by the tiny stand-ins in oracle/refshim/shims/ and permissive stubs for dace/xarray/...
"""
import importlib.abc
import importlib.machinery
import os
import sys
import types

REF = os.environ.get("PACE_REFERENCE", "/root/reference")
HERE = os.path.dirname(os.path.abspath(__file__))

# L79 hybrid coefficients: literals from the reference's own notebook
# examples/notebooks/generate_eta_file_netcdf.ipynb cells 6-7 (eta79.nc itself is not vendored).
AK79 = [
    3.000000e+02, 6.467159e+02, 1.045222e+03, 1.469188e+03, 1.897829e+03,
    2.325385e+03, 2.754396e+03, 3.191294e+03, 3.648332e+03, 4.135675e+03,
    4.668282e+03, 5.247940e+03, 5.876271e+03, 6.554716e+03, 7.284521e+03,
    8.066738e+03, 8.902188e+03, 9.791482e+03, 1.073499e+04, 1.162625e+04,
    1.237212e+04, 1.299041e+04, 1.349629e+04, 1.390277e+04, 1.422098e+04,
    1.446058e+04, 1.462993e+04, 1.473633e+04, 1.478617e+04, 1.478511e+04,
    1.473812e+04, 1.464966e+04, 1.452370e+04, 1.436382e+04, 1.417324e+04,
    1.395491e+04, 1.371148e+04, 1.344540e+04, 1.315890e+04, 1.285407e+04,
    1.253280e+04, 1.219685e+04, 1.184788e+04, 1.148739e+04, 1.111682e+04,
    1.073748e+04, 1.035062e+04, 9.957395e+03, 9.558875e+03, 9.156069e+03,
    8.749922e+03, 8.341315e+03, 7.931065e+03, 7.519942e+03, 7.108648e+03,
    6.698281e+03, 6.290007e+03, 5.884984e+03, 5.484372e+03, 5.089319e+03,
    4.700960e+03, 4.320421e+03, 3.948807e+03, 3.587201e+03, 3.236666e+03,
    2.898237e+03, 2.572912e+03, 2.261667e+03, 1.965424e+03, 1.685079e+03,
    1.421479e+03, 1.175419e+03, 9.476516e+02, 7.388688e+02, 5.497130e+02,
    3.807626e+02, 2.325417e+02, 1.054810e+02, -8.381903e-04, 0.000000e+00,
]
BK79 = [
    0., 0., 0., 0., 0., 0.,
    0., 0., 0., 0., 0., 0.,
    0., 0., 0., 0., 0., 0.,
    0., 0.00106595, 0.00412866, 0.00900663, 0.01554263, 0.02359921,
    0.03305481, 0.0438012, 0.05574095, 0.06878554, 0.08285347, 0.09786981,
    0.1137643, 0.130471, 0.1479275, 0.1660746, 0.1848558, 0.2042166,
    0.2241053, 0.2444716, 0.2652672, 0.286445, 0.3079604, 0.3297701,
    0.351832, 0.3741062, 0.3965532, 0.4191364, 0.4418194, 0.4645682,
    0.48735, 0.5101338, 0.5328897, 0.5555894, 0.5782067, 0.6007158,
    0.6230936, 0.6452944, 0.6672683, 0.6889648, 0.7103333, 0.7313231,
    0.7518838, 0.7719651, 0.7915173, 0.8104913, 0.828839, 0.846513,
    0.8634676, 0.8796583, 0.8950421, 0.9095779, 0.9232264, 0.9359506,
    0.9477157, 0.9584892, 0.9682413, 0.9769447, 0.9845753, 0.9911126,
    0.9965372, 1.,
]


NPY_DISABLE = "AVX512F AVX512CD AVX512_SKX AVX512_CLX AVX512_CNL AVX512_ICL AVX512_SPR"


class _StubMeta(type):
    """Classes whose every attribute is another stub class; supports X[...] and X(...)."""

    def __getattr__(cls, name):
        if name.startswith("__") and name.endswith("__"):
            raise AttributeError(name)
        return _StubMeta(name, (), {"_is_stub": True})

    def __getitem__(cls, item):
        return cls

    def __call__(cls, *a, **k):
        if "_is_stub" not in cls.__dict__:  # a real class deriving from a stub base
            return type.__call__(cls, *a, **k)
        if len(a) == 1 and callable(a[0]) and not k and not isinstance(a[0], type):
            return a[0]  # used as a bare decorator
        return type.__call__(cls)

    def __or__(cls, other):
        return cls

    def __ror__(cls, other):
        return cls


class _StubModule(types.ModuleType):
    def __getattr__(self, name):
        if name.startswith("__") and name.endswith("__"):
            raise AttributeError(name)
        v = _StubMeta(name, (), {"_is_stub": True})
        setattr(self, name, v)
        return v


class _StubFinder(importlib.abc.MetaPathFinder, importlib.abc.Loader):
    ROOTS = ("dace", "xarray", "f90nml", "cftime", "zarr", "netCDF4", "dacite")

    def find_spec(self, fullname, path=None, target=None):
        if fullname.split(".")[0] in self.ROOTS:
            return importlib.machinery.ModuleSpec(fullname, self, is_package=True)
        return None

    def create_module(self, spec):
        m = _StubModule(spec.name)
        m.__path__ = []
        return m

    def exec_module(self, module):
        pass


_done = False


def bootstrap(gt_cache_root=None):
    """Make `import ndsl, pyFV3` work against the read-only reference tree."""
    global _done
    if _done:
        return
    # Pin the oracle's transcendentals to the host libm: NumPy otherwise dispatches exp/log/pow to
    # AVX512 SVML kernels whose last-bit results are CPU-dependent (hard parts).
    if "numpy" in sys.modules and os.environ.get("NPY_DISABLE_CPU_FEATURES") != NPY_DISABLE:
        raise RuntimeError("bootstrap() must run before numpy is imported (NPY_DISABLE_CPU_FEATURES)")
    os.environ["NPY_DISABLE_CPU_FEATURES"] = NPY_DISABLE
    import numpy as np

    os.environ.setdefault("PYTHONDONTWRITEBYTECODE", "1")
    sys.dont_write_bytecode = True
    os.environ.setdefault("GT_CACHE_ROOT", gt_cache_root or "/tmp/pace_b200_gtcache")
    os.environ.setdefault("GT_CACHE_DIR_NAME", ".gt_cache")
    os.makedirs(os.environ["GT_CACHE_ROOT"], exist_ok=True)
    for p in (
        os.path.join(HERE, "shims"),
        os.path.join(REF, "NDSL/external/gt4py/src"),
        os.path.join(REF, "NDSL"),
        os.path.join(REF, "pyFV3"),
        os.path.join(REF, "pySHiELD"),   # physics package: only its update/ (dycore <-> physics coupling) is used
    ):
        if p not in sys.path:
            sys.path.insert(0, p)

    # gt4py.next breaks on py3.12; the cartesian frontend does not need it
    sys.modules.setdefault("gt4py.next", types.ModuleType("gt4py.next"))
    # NumPy-2 compat aliases
    if not hasattr(np, "byte_bounds"):
        np.byte_bounds = np.lib.array_utils.byte_bounds
    if not hasattr(np, "product"):
        np.product = np.prod
    import numpy.lib._index_tricks_impl as _it

    sys.modules.setdefault("numpy.lib.index_tricks", _it)

    import gt4py  # noqa: F401
    import gt4py.cartesian.gtscript  # noqa: F401
    import gt4py.storage  # noqa: F401

    sys.meta_path.append(_StubFinder())
    import dace
    import dace.frontend.python.interface as _dfi

    _dfi.nounroll = lambda x: x
    dace.nounroll = lambda x: x

    class _Cfg:
        _cfg_filename = None

        @staticmethod
        def set(*a, **k):
            pass

        @staticmethod
        def get(*a, **k):
            return None

        @staticmethod
        def get_bool(*a, **k):
            return False

    import dace.config as _dc

    _dc.Config = _Cfg
    dace.Config = _Cfg

    import ndsl.grid.eta as _eta

    def _coeffs(km, eta_file=None):
        if km != 79:
            raise ValueError("only the in-repo L79 table is available (km=%d)" % km)
        ak = np.array(AK79, dtype=np.float64)
        bk = np.array(BK79, dtype=np.float64)
        return _eta.HybridPressureCoefficients(18, ak[0], ak, bk)

    _eta.set_hybrid_pressure_coefficients = _coeffs
    import ndsl.grid.generation as _gen

    if hasattr(_gen, "set_hybrid_pressure_coefficients"):
        _gen.set_hybrid_pressure_coefficients = _coeffs
    _done = True
