"""SQ(x) == Python's x ** 2 (libm pow(x, 2.0)): host build of csrc/pow2_exact.cuh against libm
(CPU), device build against the host's libm (GPU)."""
import os
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_host_emulation_matches_libm(tmp_path):
    exe = tmp_path / "pow2_check"
    subprocess.run(["g++", "-O2", "-ffp-contract=off", "-fno-builtin-pow", "-mfma", "-o", str(exe),
                    os.path.join(ROOT, "tests", "native", "pow2_check.cpp"), "-lm"], check=True)
    out = subprocess.run([str(exe), "20000000", "5"], capture_output=True, text=True)
    assert out.returncode == 0, out.stdout
    assert "table-path mismatches: 0  pow2_exact mismatches: 0" in out.stdout
    # the emulation is only needed because pow(x,2) != x*x for ~8e-4 of inputs on this libm
    assert "pow!=x*x: 0 " not in out.stdout


def test_tables_match_this_libm(tmp_path):
    """pow2_tables.h is generated from the image's libm.so.6; regenerate and compare."""
    libm = "/lib/x86_64-linux-gnu/libm.so.6"
    if not os.path.exists(libm):
        pytest.skip("no glibc libm at the expected path")
    committed = open(os.path.join(ROOT, "nclone_b200", "csrc", "pow2_tables.h")).read()
    src = open(os.path.join(ROOT, "tools", "extract_pow_tables.py")).read().replace(
        'os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "nclone_b200", "csrc", "pow2_tables.h")',
        repr(str(tmp_path / "t.h")))
    (tmp_path / "x.py").write_text(src)
    subprocess.run(["python", str(tmp_path / "x.py")], check=True, capture_output=True)
    assert (tmp_path / "t.h").read_text().split("\n", 1)[1] == committed.split("\n", 1)[1]


@pytest.mark.gpu
def test_device_sq_matches_python_pow():
    import ctypes as C
    import torch
    from nclone_b200._lib import lib, check
    L = lib()
    assert L.nppsim_sq_mode() == 1
    rng = np.random.default_rng(3)
    n = 1 << 22
    parts = [rng.uniform(-1100, 1100, n), rng.uniform(-12, 12, n), rng.uniform(-1, 1, n),
             np.ldexp(rng.uniform(1, 2, n), rng.integers(-40, 20, n)) * rng.choice([-1.0, 1.0], n),
             np.array([0.0, -0.0, 1.0, -1.0, 2.0, 12.0, 24.0, 5.0, 29.0, 19.0, 1e-300, 1e300, np.inf, 0.5, 3.0, 1e-160, 1e160])]
    x = np.concatenate(parts)
    want = np.array([float(v) ** 2 if abs(v) < 1e150 else v * v for v in x[-17:]])
    xd = torch.from_numpy(x).cuda(); od = torch.empty_like(xd)
    check(L.nppsim_debug_sq(C.c_void_p(xd.data_ptr()), C.c_void_p(od.data_ptr()), len(x), None))
    got = od.cpu().numpy()
    libm = C.CDLL("libm.so.6"); libm.pow.restype = C.c_double; libm.pow.argtypes = [C.c_double, C.c_double]
    # vectorised reference: numpy's float_power goes through the same libm pow for float64
    ref = np.float_power(x[:-17], 2.0)
    spot = rng.integers(0, len(x) - 17, 2000)
    for k in spot:
        assert ref[k] == libm.pow(float(x[k]), 2.0) == float(x[k]) ** 2
    bad = np.nonzero(got[:-17].view(np.int64) != ref.view(np.int64))[0]
    assert bad.size == 0, (x[bad[:5]], got[bad[:5]], ref[bad[:5]])
    assert np.array_equal(got[-17:].view(np.int64), want.view(np.int64))
    assert np.count_nonzero(ref != x[:-17] * x[:-17]) > 1000      # the emulation is doing real work


@pytest.mark.gpu
def test_device_div2_is_ieee_division():
    """csrc/geom.cuh div2 (two quotients by one denominator off a single reciprocal refinement) must return the bits of
    the device's own IEEE `/` for every input: 2^28 random triples over the kernels' working range (24*dx / dist),
    over wide exponent ranges, and over the special values that must take the guarded fallback."""
    import ctypes as C
    import torch
    from nclone_b200._lib import lib, check
    L = lib()
    g = torch.Generator(device="cuda"); g.manual_seed(5)
    bad = torch.zeros(1, dtype=torch.int64, device="cuda")

    def run(a1, a2, b):
        check(L.nppsim_debug_div2(C.c_void_p(a1.data_ptr()), C.c_void_p(a2.data_ptr()), C.c_void_p(b.data_ptr()),
                                  C.c_void_p(bad.data_ptr()), a1.numel(), None))
    n = 1 << 24
    for rep in range(8):
        dx = (torch.rand(n, generator=g, device="cuda", dtype=torch.float64) - 0.5) * 96
        dy = (torch.rand(n, generator=g, device="cuda", dtype=torch.float64) - 0.5) * 96
        run(24.0 * dx, 24.0 * dy, torch.sqrt(dx * dx + dy * dy))
    for rep in range(8):
        e = lambda: torch.exp2((torch.rand(n, generator=g, device="cuda", dtype=torch.float64) - 0.5) * 600) * \
            torch.where(torch.rand(n, generator=g, device="cuda") < 0.5, -1.0, 1.0).double() * (1 + torch.rand(n, generator=g, device="cuda", dtype=torch.float64))
        run(e(), e(), e())
    sp = torch.tensor([0.0, -0.0, 1.0, -1.0, 5e-324, 2.2e-308, 1e-300, 1e300, 1.7e308, float("inf"), -float("inf"), float("nan"), 3.0, 1e-160, 1e160],
                      dtype=torch.float64, device="cuda")
    k = sp.numel()
    a1 = sp.repeat_interleave(k * k); a2 = sp.repeat(k).repeat_interleave(k); b = sp.repeat(k * k)
    run(a1.contiguous(), a2.contiguous(), b.contiguous())
    torch.cuda.synchronize()
    assert int(bad.item()) == 0
