"""Stage G3 on its own: the pruned 2-D transform of csrc/pfft.cu (row pass -- cuFFT's or, CIP_FFT=own, the own one --
two own column passes, crop in image order) against numpy's FFT on the same plane.  Unnormalised, exp(+i) sign -- what cuFFT's CUFFT_INVERSE gives and
what the w-stacking algorithm of SURVEY.md appendix A needs.  Tolerances: float32 transforms of n points carry a
relative rms error of a few 1e-7 (checked at 2e-6 of the output rms), float64 ones 1e-15 (checked at 1e-13)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def cabi(built_lib):
    import torch
    assert torch.cuda.is_available(), "gpu tests need a CUDA device"
    from ska_sdp_cip import _cabi
    return _cabi


def reference_crop(plane, nx, ny):
    nu, nv = plane.shape
    full = np.fft.ifft2(plane.astype(np.complex128)) * (nu * nv)
    iu = (np.arange(nx) - nx // 2) % nu
    iv = (np.arange(ny) - ny // 2) % nv
    return full[np.ix_(iu, iv)]


# every radix (16, 8, 4, 2, 3, 5, 7), one- and many-stage sizes, n1 = 1 (nu <= 128), ny not a multiple of 16,
# no crop at all, and the shapes of BASELINE configs[0] (sigma 1.5 and 2)
SIZES = [(16, 16, 16, 16), (32, 64, 16, 32), (96, 128, 64, 80), (98, 150, 50, 74), (210, 420, 128, 200),
         (384, 640, 256, 400), (1000, 1536, 512, 1024), (1536, 1536, 1024, 1024), (2048, 2048, 1024, 1024),
         (686, 2 * 3 * 5 * 7 * 4, 300, 422), (1280, 2560, 1280, 2560)]


@pytest.fixture(params=["default", "own"])
def row_pass(request):
    """default: cuFFT does the contiguous row pass (float32, image width a multiple of 4); own: rows_kernel_p."""
    import os
    if request.param == "own":
        os.environ["CIP_FFT"] = "own"
    yield request.param
    os.environ.pop("CIP_FFT", None)


@pytest.mark.parametrize("nu,nv,nx,ny", SIZES)
@pytest.mark.parametrize("dtype,tol", [(np.complex64, 2e-6), (np.complex128, 1e-13)])
def test_pruned_transform_matches_numpy(cabi, row_pass, nu, nv, nx, ny, dtype, tol):
    import torch
    if dtype == np.complex128 and row_pass == "default":
        pytest.skip("float64 planes always take the own row pass")
    rng = np.random.default_rng(nu * 131 + nv)
    plane = (rng.standard_normal((nu, nv)) + 1j * rng.standard_normal((nu, nv))).astype(dtype)
    ctx = cabi.Context.get(-1)
    out = cabi.plane_fft_crop(ctx, torch.as_tensor(plane).cuda(), nx, ny).cpu().numpy()
    ref = reference_crop(plane, nx, ny)
    assert out.shape == (nx, ny)
    err = np.sqrt(np.mean(np.abs(out - ref) ** 2) / np.mean(np.abs(ref) ** 2))
    assert err <= tol, err


def test_single_tone_lands_on_its_pixel(cabi, row_pass):
    """A plane that is one complex exponential transforms to a single retained cell (indexing and sign)."""
    import torch
    nu, nv, nx, ny = 768, 640, 512, 320
    fu, fv = -37, 101  # frequencies inside the retained window
    u = np.arange(nu)[:, None]
    v = np.arange(nv)[None, :]
    plane = np.exp(-2j * np.pi * (u * fu / nu + v * fv / nv)).astype(np.complex64)
    out = cabi.plane_fft_crop(cabi.Context.get(-1), torch.as_tensor(plane).cuda(), nx, ny).cpu().numpy()
    peak = np.unravel_index(np.argmax(np.abs(out)), out.shape)
    assert peak == (fu + nx // 2, fv + ny // 2)
    assert abs(out[peak] - nu * nv) <= 1e-3 * nu * nv
    out[peak] = 0
    assert np.abs(out).max() <= 2e-4 * nu * nv


def test_headline_plane_size(cabi, row_pass):
    """The 6144^2 plane / 4096^2 image of BASELINE configs[1] (float32)."""
    import torch
    nu = nv = 6144
    nx = ny = 4096
    g = torch.Generator(device="cuda").manual_seed(5)
    plane = torch.complex(torch.randn((nu, nv), device="cuda", generator=g), torch.randn((nu, nv), device="cuda", generator=g))
    out = cabi.plane_fft_crop(cabi.Context.get(-1), plane, nx, ny)
    full = torch.fft.ifft2(plane.to(torch.complex128)) * (nu * nv)
    iu = (torch.arange(nx, device="cuda") - nx // 2) % nu
    iv = (torch.arange(ny, device="cuda") - ny // 2) % nv
    ref = full[iu][:, iv]
    err = float(torch.sqrt(torch.mean(torch.abs(out - ref) ** 2) / torch.mean(torch.abs(ref) ** 2)))
    assert err <= 2e-6, err
