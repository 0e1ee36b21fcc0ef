"""CPU tests of the drop-in boundary: the C-ABI library loads, exports every symbol that
include/miraculix_b200.h declares, and keeps the reference's error conventions.  No compute
call needs a GPU here."""
import os
import re
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "miraculix_b200.h")


def declared_symbols():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    src = re.sub(r"typedef\s+[^;{]*\(\s*\*[^;]*;", "", src)  # function-pointer typedefs
    return sorted(set(re.findall(r"\b([A-Za-z_][A-Za-z0-9_]*)\s*\([^;{]*\)\s*;", src)))


def test_header_declares_reference_surface():
    names = declared_symbols()
    for ref_name in ("setOptions_compressed", "plink2compressed", "dgemm_compressed",
                     "free_compressed", "get_compressed_freq", "snp_multiply_gpu"):
        assert ref_name in names


def test_library_exports_every_declared_symbol(lib):
    from miraculix_b200 import api
    names = declared_symbols()
    assert len(names) >= 20
    for n in names:
        assert hasattr(lib, n), f"{n} declared in the header but not exported"
        assert n in api.SIGNATURES, f"{n} has no ctypes signature in api.SIGNATURES"


def _run(code: str):
    env = dict(os.environ, PYTHONPATH=ROOT)
    return subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, env=env,
                          timeout=300)


def test_bad_trans_exits_99(lib):
    """5codesAPI.c:73-77: anything but N/n/T/t/Y/y -> exit(99)."""
    r = _run("from miraculix_b200 import api; import ctypes as C; L = api.load_library();"
             "L.dgemm_compressed(b'X', None, 1, None, 1, None, 1)")
    assert r.returncode == 99


def test_inadmissible_gpu_options_exit_failure(lib):
    """5codesChar.cc:192-193: gpu needs external freq, ignore_missings, no normalising."""
    for flags in ("1,0,0,0,1,0,0,1,0,0",   # use_miraculix_freq = 1
                  "1,0,0,0,0,0,0,0,0,0",   # ignore_missings = 0
                  "1,0,0,0,1,0,1,0,0,0"):  # do_normalize = 1
        r = _run(f"from miraculix_b200 import api; api.load_library().setOptions_compressed({flags})")
        assert r.returncode == 1, flags
        assert "Fortran frequency" in r.stderr
    r = _run("from miraculix_b200 import api; api.set_options(use_gpu=True, not_center=True)")
    assert r.returncode == 0


def test_invalid_handle_is_reported_not_crashed(lib):
    from miraculix_b200 import api
    buf = np.zeros(4)
    lib.dgemm_compressed(b"N", None, 1, buf.ctypes.data, 4, buf.ctypes.data, 4)
    assert lib.mb200_last_status() != 0
    assert b"invalid" in lib.mb200_last_message()
    lib.free_compressed(None)  # NULL is a no-op


def test_host_argument_checks():
    from miraculix_b200 import api
    with pytest.raises(ValueError):
        api.init_compressed(np.zeros((3, 3), np.uint8), 3, 8, np.zeros(3), 1)  # wrong bed shape
    with pytest.raises(ValueError):
        api.grm(np.zeros((4, 2), np.uint8), 8, 4, do_scale=True, allele_freq=np.zeros(3))
    with pytest.raises(ValueError):
        api.snp_crossprod(np.zeros((4, 3), np.uint8), 8, 4, is_snpmajor=False)


def test_no_device_fails_loudly(lib):
    """On a box without a GPU the product must error, never fall back to a CPU path."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    from miraculix_b200 import api
    plink = np.zeros((4, 1), np.uint8)
    with pytest.raises(api.MiraculixError):
        api.init_compressed(plink, 4, 4, np.zeros(4), 1)
    with pytest.raises(api.MiraculixError):
        api.snp_crossprod(np.zeros((4, 1), np.uint8), 4, 4, is_snpmajor=False)


def test_product_never_touches_the_oracle(lib):
    """The oracle is test infrastructure: nothing under miraculix_b200/ imports, opens or links it,
    and the shipped library depends on neither the oracle's nor the reference's shared object."""
    pkg = os.path.join(ROOT, "miraculix_b200")
    pat = re.compile(r"^\s*(from|import)\s+oracle\b|oracle/|oracle\.|_ref\b|ref_miraculix", re.M)
    for base, _, files in os.walk(pkg):
        for name in files:
            if not name.endswith((".py", ".cu", ".cuh", ".h", ".cpp")):
                continue
            src = open(os.path.join(base, name)).read()
            src = re.sub(r'""".*?"""', "", src, flags=re.S)  # docstrings may mention the tests
            assert not pat.search(src), f"{name} refers to the oracle"
    from miraculix_b200 import api
    out = subprocess.run(["ldd", api.LIB_PATH], capture_output=True, text=True).stdout
    assert "oracle" not in out and "ref_miraculix" not in out
    mods = _run("import sys, miraculix_b200, miraculix_b200.api, miraculix_b200.sharded;"
                "print([m for m in sys.modules if m.split('.')[0] == 'oracle'])")
    assert mods.returncode == 0 and mods.stdout.strip() == "[]", mods.stdout + mods.stderr


def test_device_list_without_a_gpu_fails_loudly(lib):
    """MIRACULIX_NGPUS / CUDA_DEVICE lists / mb200_set_devices select the sharded path; without a CUDA
    device every entry point refuses with a message and a NULL handle (no CPU path, no crash).  Runs
    in a subprocess with the GPUs hidden, so it means the same thing on a GPU box."""
    code = (
        "import os, numpy as np, ctypes as C\n"
        "os.environ['CUDA_VISIBLE_DEVICES'] = ''\n"
        "from miraculix_b200 import api\n"
        "L = api.load_library()\n"
        "plink = np.zeros((600, 3), dtype=np.uint8); f = np.zeros(600); h = C.c_void_p(None)\n"
        "for how in ('env', 'list', 'api'):\n"
        "    os.environ.pop('MIRACULIX_NGPUS', None); os.environ.pop('CUDA_DEVICE', None); api.set_devices(None)\n"
        "    if how == 'env': os.environ['MIRACULIX_NGPUS'] = '2'\n"
        "    if how == 'list': os.environ['CUDA_DEVICE'] = '0,1'\n"
        "    if how == 'api': api.set_devices([0, 1])\n"
        "    L.plink2compressed(plink.ctypes.data, plink.ctypes.data, 600, 12, f.ctypes.data, 1, C.byref(h))\n"
        "    assert not h.value and L.mb200_last_status() != 0, how\n"
        "    ans = np.zeros((12, 12))\n"
        "    assert L.snp_multiply_gpu(plink.ctypes.data, 12, 12, ans.ctypes.data, False) == 1\n"
        "assert L.mb200_shard_count(None) == -1\n"
        "print('refused')\n")
    r = _run(code)
    assert r.returncode == 0 and "refused" in r.stdout, r.stdout + r.stderr
    assert "no usable CUDA device" in r.stderr


def test_copy_pool_serves_concurrent_callers(lib):
    """The staging copy pool (csrc/stager.cu): eight threads copy 2-D blocks of different shapes at
    the same time, many times over, sharing the helper threads -- every destination must equal its
    source, bytes outside the copied columns must stay untouched (host logic, no GPU)."""
    import threading

    import numpy as np
    errors = []

    def worker(seed):
        rng = np.random.default_rng(seed)
        for it in range(12):
            width = int(rng.integers(1, 3_000_000)) if it % 3 else int(rng.integers(1, 5000))
            height = int(rng.integers(1, 9)) if width > 100_000 else int(rng.integers(1, 700))
            spitch, dpitch = width + int(rng.integers(0, 64)), width + int(rng.integers(0, 64))
            src = rng.integers(0, 256, size=spitch * height, dtype=np.uint8)
            dst = np.full(dpitch * height, 0xA5, dtype=np.uint8)
            rc = lib.mb200_debug_parallel_copy(dst.ctypes.data, dpitch, src.ctypes.data, spitch, width, height)
            s2, d2 = src.reshape(height, spitch), dst.reshape(height, dpitch)
            if rc != 0 or not np.array_equal(d2[:, :width], s2[:, :width]) or not np.all(d2[:, width:] == 0xA5):
                errors.append((seed, it, width, height, rc))

    threads = [threading.Thread(target=worker, args=(s,)) for s in range(8)]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    assert not errors, errors[:3]
    big = np.arange(64 << 20, dtype=np.uint8)          # one large flat copy: cut into many pieces
    out = np.zeros_like(big)
    assert lib.mb200_debug_parallel_copy(out.ctypes.data, big.size, big.ctypes.data, big.size, big.size, 1) == 0
    assert np.array_equal(out, big)
    assert lib.mb200_debug_parallel_copy(None, 8, big.ctypes.data, 8, 8, 1) != 0


def test_c_demo_compiles_as_c99_and_fails_loudly_without_a_gpu(lib):
    """examples/c_abi_demo.c: the reference's call sequence in plain C.  The header must be valid
    C99 (the bindings are C / Fortran / Julia), the program must link against the in-tree library,
    and on a box without a GPU it must report the missing device instead of computing anything."""
    import shutil
    import subprocess
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    gcc = shutil.which("gcc")
    if not gcc:
        pytest.skip("no gcc")
    out = os.path.join(root, "miraculix_b200", "build", "c_abi_demo")
    os.makedirs(os.path.dirname(out), exist_ok=True)
    libdir = os.path.join(root, "miraculix_b200")
    r = subprocess.run([gcc, "-std=c99", "-Wall", "-Wextra", "-Werror", "-O2", "-I", os.path.join(root, "include"),
                        os.path.join(root, "examples", "c_abi_demo.c"), "-o", out, "-L", libdir,
                        "-lmiraculix_b200", f"-Wl,-rpath,{libdir}", "-lm"], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    try:
        import torch
        has_gpu = torch.cuda.is_available()
    except Exception:
        has_gpu = False
    run = subprocess.run([out], capture_output=True, text=True, timeout=300)
    if has_gpu:
        assert run.returncode == 0 and "ok" in run.stdout, run.stdout + run.stderr
    else:
        assert run.returncode == 2 and "no handle" in run.stdout and "no CPU path" in run.stdout, run.stdout
