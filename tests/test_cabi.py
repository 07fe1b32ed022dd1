"""The C-ABI library without a GPU: it loads, exports every symbol include/cip.h declares, its
host-only entry points (planning, kernel table, kernel Fourier transform) agree with the oracle,
and the device entry points fail loudly instead of falling back to the CPU."""
import ctypes as C
import re
from pathlib import Path

import numpy as np
import pytest

from oracle import plan as oplan, wstack

ROOT = Path(__file__).resolve().parent.parent
HEADER = ROOT / "include" / "cip.h"


def declared_symbols():
    text = re.sub(r"/\*.*?\*/", "", HEADER.read_text(), flags=re.S)
    return sorted(set(re.findall(r"\b(cip_[a-z0-9_]+)\s*\(", text)))


def test_header_declares_the_documented_entry_points():
    names = declared_symbols()
    for must in ("cip_invert", "cip_tile_bin", "cip_stokes_i", "cip_grid_accumulate", "cip_correct_image"):
        assert must in names


def test_library_exports_every_declared_symbol(built_lib):
    from ska_sdp_cip import _cabi
    names = declared_symbols()
    assert names, "no declarations parsed from include/cip.h"
    for name in names:
        assert hasattr(built_lib, name), f"{name} declared in include/cip.h but not exported"
    assert sorted(_cabi.EXPORTS) == names
    assert built_lib.cip_version().decode().startswith("cip_b200")


def test_struct_layouts_match_header_field_order():
    """The ctypes mirrors must list the header's fields in the header's order."""
    from ska_sdp_cip import _cabi
    text = re.sub(r"/\*.*?\*/", "", HEADER.read_text(), flags=re.S)
    for cname, cls in (("cip_invert_args", _cabi.InvertArgs), ("cip_scan_result", _cabi.ScanResult),
                       ("cip_plan", _cabi.Plan), ("cip_report", _cabi.Report)):
        body = re.search(r"typedef struct \{([^{}]*)\}\s*" + cname + r"\s*;", text, flags=re.S).group(1)
        fields = []
        for decl in body.split(";"):
            decl = decl.strip()
            if not decl:
                continue
            fields += [re.sub(r"[\s\*]", "", n.split()[-1]) for n in decl.split(",")]
        assert fields == [n for n, _ in cls._fields_], cname


PLAN_CASES = [
    # nx, px, eps, wmin, wmax, nvis, fp64, wstack
    (1024, 1e-5, 1e-4, 0.0, 1000.0, 10 ** 6, False, True),
    (4096, 7.3e-7, 1e-4, 0.0, 3.3e5, 168689664, False, True),
    (8192, 1.34e-5, 1e-4, 0.0, 1.2e4, 535822336, False, True),
    (512, 2e-4, 1e-5, 10.0, 2.7e4, 10 ** 7, True, True),
    (256, 1e-4, 1e-3, 0.0, 50.0, 1000, False, True),
    (256, 1e-4, 1e-4, 0.0, 50.0, 1000, False, False),
    (2048, 2.4e-5, 1e-6, 0.0, 900.0, 3 * 10 ** 5, True, True),
]


@pytest.mark.parametrize("nx,px,eps,wmin,wmax,nvis,fp64,wst", PLAN_CASES)
def test_plan_matches_oracle(built_lib, nx, px, eps, wmin, wmax, nvis, fp64, wst):
    from ska_sdp_cip import _cabi
    p = _cabi.plan_make(nx, nx, px, px, eps, wmin, wmax, nvis, do_wstacking=wst, fp64=fp64)
    o = oplan.make_plan(nx, nx, px, px, eps, wmin, wmax, nvis, fp64=fp64, do_wstacking=wst)
    assert (p.sigma, p.support, p.nu, p.nv, p.nplanes) == (o.sigma, o.W, o.nu, o.nv, o.nplanes)
    assert p.beta == pytest.approx(o.beta, rel=1e-12)
    assert p.dw == pytest.approx(o.dw, rel=1e-12)
    assert p.w0 == pytest.approx(o.w0, rel=1e-12, abs=1e-9)
    assert p.nshift == pytest.approx(o.nshift, rel=1e-12, abs=1e-300)


def test_plan_overrides_and_errors(built_lib):
    from ska_sdp_cip import _cabi
    p = _cabi.plan_make(512, 512, 1e-5, 1e-5, 1e-4, 0.0, 100.0, 1000, sigma=2.0, support=6)
    assert (p.sigma, p.support) == (2.0, 6)
    with pytest.raises(ValueError):
        _cabi.plan_make(511, 512, 1e-5, 1e-5, 1e-4, 0.0, 100.0, 1000)        # odd image size
    with pytest.raises(ValueError):
        _cabi.plan_make(512, 512, 1e-5, 1e-5, 1e-30, 0.0, 100.0, 1000)       # unreachable accuracy
    with pytest.raises(ValueError):
        _cabi.plan_make(512, 512, -1e-5, 1e-5, 1e-4, 0.0, 100.0, 1000)       # negative pixel size


def test_kernel_table_matches_design_table(built_lib):
    for sigma, tab in oplan.ES_TABLE.items():
        for W, (beta, err) in tab.items():
            b, e = C.c_double(), C.c_double()
            assert built_lib.cip_kernel_table(sigma, W, C.byref(b), C.byref(e)) == 0
            assert b.value == pytest.approx(beta, rel=1e-12) and e.value == pytest.approx(err, rel=1e-12)
    b, e = C.c_double(), C.c_double()
    assert built_lib.cip_kernel_table(3.0, 6, C.byref(b), C.byref(e)) == 1  # CIP_EINVAL


@pytest.mark.parametrize("sigma,W", [(1.5, 7), (2.0, 6), (2.0, 12), (1.25, 9)])
def test_kernel_ft_matches_oracle_quadrature(built_lib, sigma, W):
    from ska_sdp_cip import _cabi
    p = _cabi.plan_make(256, 256, 1e-4, 1e-4, 1e-4, 0.0, 50.0, 1000, sigma=sigma, support=W)
    o = oplan.make_plan(256, 256, 1e-4, 1e-4, 1e-4, 0.0, 50.0, 1000, sigma=sigma, W=W)
    xi = np.linspace(-0.5 / sigma, 0.5 / sigma, 301)
    got = _cabi.kernel_ft(p, xi)
    want = wstack.kernel_ft(o, xi)
    assert np.allclose(got, want, rtol=1e-11, atol=0)
    assert got.min() > 0 and got[150] == got.max()


def test_no_cpu_fallback_without_a_device(built_lib):
    """On a machine without a GPU the product must raise, never compute on the host."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("a CUDA device is present")
    from ska_sdp_cip import _cabi, wgridder
    from ska_sdp_cip.uvw_tiling import create_uvw_tile_mapping
    with pytest.raises(_cabi.CipError):
        _cabi.Context(0)
    uvw = np.zeros((4, 3))
    freq = np.array([1e9, 1.1e9])
    vis = np.ones((4, 2), np.complex64)
    with pytest.raises(_cabi.CipError):
        wgridder.ms2dirty(uvw, freq, vis, None, 64, 64, 1e-5, 1e-5, epsilon=1e-4, do_wstacking=True)
    with pytest.raises(_cabi.CipError):
        create_uvw_tile_mapping(uvw, (1.0, 1.0, 1.0), freq)


def test_missing_library_is_an_import_error(monkeypatch, tmp_path):
    from ska_sdp_cip import _cabi
    monkeypatch.setenv("CIP_B200_LIB", str(tmp_path / "nope.so"))
    monkeypatch.setattr(_cabi, "_lib", None)
    with pytest.raises(ImportError, match="no CPU fallback"):
        _cabi.lib()


def test_product_never_imports_the_oracle():
    """oracle/ is test infrastructure: nothing under the package may import it."""
    pkg = ROOT / "ska-sdp-continuum-imaging-pipeline_b200"
    for path in pkg.rglob("*"):
        if path.suffix in (".py", ".cu", ".cuh", ".h", ".cpp") and "build" not in path.parts:
            text = path.read_text()
            assert not re.search(r"^\s*(from|import)\s+oracle\b", text, flags=re.M), path
            assert "oracle/" not in text or path.suffix == ".py" and "test infrastructure" in text, path
