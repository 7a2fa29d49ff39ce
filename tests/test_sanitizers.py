"""ThreadSanitizer over the C++ control plane / ring all-reduce: 4 ranks as threads of one instrumented process
(csrc/runtime_tests/tsan_driver.cc). SURVEY 5.2."""
import os
import shutil
import subprocess

import pytest

from _mp import free_ports

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_runtime_is_clean_under_thread_sanitizer(tmp_path):
    if shutil.which("g++") is None:
        pytest.skip("no g++")
    exe = str(tmp_path / "tsan_driver")
    srcs = [os.path.join(ROOT, "csrc", "runtime", f) for f in sorted(os.listdir(os.path.join(ROOT, "csrc", "runtime")))
            if f.endswith(".cc")] + [os.path.join(ROOT, "csrc", "runtime_tests", "tsan_driver.cc")]
    b = subprocess.run(["g++", "-std=c++17", "-g", "-O1", "-fsanitize=thread", "-pthread", *srcs, "-o", exe],
                       capture_output=True, text=True)
    if b.returncode != 0 and "tsan" in (b.stderr or "").lower():
        pytest.skip("libtsan not available")
    assert b.returncode == 0, b.stderr[-2000:]
    base = free_ports(1)[0]
    # 64 KB: the driver's 400 KB all-reduces take the helper-thread path of the ring (send half on a second thread)
    env = dict(os.environ, TSAN_OPTIONS="halt_on_error=1 exitcode=66", TFDL_RING_THREAD_BYTES="65536")
    r = subprocess.run([exe, "4", str(base), "5", str(tmp_path)], capture_output=True, text=True, timeout=300, env=env)
    assert r.returncode == 0 and "ThreadSanitizer" not in r.stderr, (r.stdout[-500:], r.stderr[-3000:])
    assert "-> ok" in r.stdout


def test_runtime_is_clean_under_address_and_ub_sanitizers(tmp_path):
    """The same in-process multi-rank driver (collectives, worker pool, blocking and background gathers) under
    AddressSanitizer + UndefinedBehaviorSanitizer."""
    if shutil.which("g++") is None:
        pytest.skip("no g++")
    exe = str(tmp_path / "asan_driver")
    srcs = [os.path.join(ROOT, "csrc", "runtime", f) for f in sorted(os.listdir(os.path.join(ROOT, "csrc", "runtime")))
            if f.endswith(".cc")] + [os.path.join(ROOT, "csrc", "runtime_tests", "tsan_driver.cc")]
    b = subprocess.run(["g++", "-std=c++17", "-g", "-O1", "-fsanitize=address,undefined", "-fno-omit-frame-pointer", "-pthread",
                        *srcs, "-o", exe], capture_output=True, text=True)
    if b.returncode != 0 and ("asan" in (b.stderr or "").lower() or "ubsan" in (b.stderr or "").lower()):
        pytest.skip("libasan / libubsan not available")
    assert b.returncode == 0, b.stderr[-2000:]
    base = free_ports(1)[0]
    env = dict(os.environ, ASAN_OPTIONS="detect_leaks=0", UBSAN_OPTIONS="halt_on_error=1:print_stacktrace=1",
               TFDL_RING_THREAD_BYTES="65536")
    env.pop("LD_PRELOAD", None)
    r = subprocess.run([exe, "4", str(base), "4", str(tmp_path)], capture_output=True, text=True, timeout=300, env=env)
    assert r.returncode == 0 and "ERROR: AddressSanitizer" not in r.stderr and "runtime error" not in r.stderr, (
        r.stdout[-500:], r.stderr[-3000:])
    assert "-> ok" in r.stdout


def test_bundle_reader_survives_damaged_checkpoints(tmp_path):
    """Mutation fuzzing of the checkpoint reader under ASan + UBSan (csrc/runtime_tests/bundle_fuzz.cc): flipped bytes,
    truncation, overwritten runs and duplicated slices in the index or data file end in an error, not in a crash."""
    if shutil.which("g++") is None:
        pytest.skip("no g++")
    exe = str(tmp_path / "bundle_fuzz")
    srcs = [os.path.join(ROOT, "csrc", "runtime", f) for f in sorted(os.listdir(os.path.join(ROOT, "csrc", "runtime")))
            if f.endswith(".cc")] + [os.path.join(ROOT, "csrc", "runtime_tests", "bundle_fuzz.cc")]
    b = subprocess.run(["g++", "-std=c++17", "-g", "-O1", "-fsanitize=address,undefined", "-fno-omit-frame-pointer", "-pthread",
                        *srcs, "-o", exe], capture_output=True, text=True)
    if b.returncode != 0 and ("asan" in (b.stderr or "").lower() or "ubsan" in (b.stderr or "").lower()):
        pytest.skip("libasan / libubsan not available")
    assert b.returncode == 0, b.stderr[-2000:]
    env = dict(os.environ, ASAN_OPTIONS="detect_leaks=0", UBSAN_OPTIONS="halt_on_error=1:print_stacktrace=1")
    env.pop("LD_PRELOAD", None)
    r = subprocess.run([exe, str(tmp_path), "2500", "20261107"], capture_output=True, text=True, timeout=600, env=env)
    assert r.returncode == 0 and "-> ok" in r.stdout, (r.stdout[-500:], r.stderr[-3000:])
    opened = int(r.stdout.split("opened=")[1].split()[0])
    assert 0 < opened < 2500  # some damage is survivable (data file), most is refused
