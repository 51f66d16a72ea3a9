"""-m gpu, collected last: the allocation-free diffusion entry point.

Some hosts embed this library where they are not allowed to allocate GPU memory (the device has
been split into a fixed arena up front, or the call happens where the CUDA allocator must stay
untouched).  ve_diffusion_static serves them: caller-owned device buffers in, results in place,
work space taken from an array compiled into the library.  It must give exactly what the
allocating entry point ve_diffusion gives — checked here on two calls in a row, the second with
different data so that anything left in the shared work space by the first would show.
"""
import numpy as np
import pytest

from util import ROOT, bit_diff  # noqa: F401
from veloren_b200.erosion import ErosionContext, VeError, diffusion_static

pytestmark = pytest.mark.gpu


def _case(seed, x_lg, y_lg):
    rng = np.random.default_rng(seed)
    n = (1 << x_lg) * (1 << y_lg)
    h = rng.normal(size=n) * 50.0
    b = h - rng.random(n) * 3.0
    kd = 1.25e-3 * (1.0 + 20.0 * rng.random(n))
    return h, b, kd


def test_static_diffusion_matches_allocating_entry_point_twice():
    import torch
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    dt = 32000.0
    for call, (seed, x_lg, y_lg) in enumerate(((11, 8, 8), (12, 9, 7))):
        h, b, kd = _case(seed, x_lg, y_lg)
        ref = ErosionContext(x_lg, y_lg, device=0)
        h_ref, b_ref = ref.diffusion(dt, h, b, kd)            # existing entry point (allocates)
        ref.close()
        d_h = torch.from_numpy(h.copy()).to(dev)
        d_b = torch.from_numpy(b.copy()).to(dev)
        d_kd = torch.from_numpy(kd.copy()).to(dev)
        torch.cuda.synchronize()
        diffusion_static(x_lg, y_lg, dt, d_h.data_ptr(), d_b.data_ptr(), d_kd.data_ptr(),
                         torch.cuda.current_stream().cuda_stream)
        h_new, b_new = d_h.cpu().numpy(), d_b.cpu().numpy()
        # same bar as the other parity tests: bit-identical (tolerance 1e-9 relative is the contract)
        assert np.allclose(h_new, h_ref, rtol=1e-9, atol=0.0) and np.allclose(b_new, b_ref, rtol=1e-9, atol=0.0), call
        assert bit_diff(h_new, h_ref) == 0 and bit_diff(b_new, b_ref) == 0, call
        assert not np.array_equal(h_new, h), "diffusion did nothing"


def test_static_diffusion_rejects_maps_beyond_its_work_space():
    import torch
    dev = torch.device("cuda", 0)
    t = torch.zeros(16, dtype=torch.float64, device=dev)
    with pytest.raises(VeError) as ei:
        diffusion_static(10, 10, 1.0, t.data_ptr(), t.data_ptr(), t.data_ptr(), 0)   # 1024x1024 > 512x512
    assert ei.value.code == -7 and "512x512" in str(ei.value)
