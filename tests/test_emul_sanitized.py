"""The decoders' device source (tests/emul: the same .cuh files compiled for the host, one lane) under
AddressSanitizer + UndefinedBehaviorSanitizer (race detection / sanitizers row).
compute-sanitizer is closed on the GPU pool, so this is the memory-safety check of the decode logic:
every table, ring, queue and sequence-arena access of k_lane / k_big / the group decoder / the z_stream
decoder runs here with red zones around each buffer.  The quick subset below is part of the CPU tier;
scripts/emul_sanitize.sh runs every emulation test this way and writes profiles/r01_emul_asan_ubsan.txt."""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _libasan():
    try:
        p = subprocess.run(["gcc", "-print-file-name=libasan.so"], capture_output=True, text=True, check=True).stdout.strip()
    except (OSError, subprocess.CalledProcessError):
        return None
    return p if os.path.isabs(p) and os.path.exists(p) else None


def sanitized_pytest(args, timeout):
    asan = _libasan()
    if asan is None:
        pytest.skip("gcc's libasan.so not found")
    env = dict(os.environ, ZSB_EMUL_SANITIZE="1", LD_PRELOAD=asan, ASAN_OPTIONS="detect_leaks=0:abort_on_error=0",
               UBSAN_OPTIONS="print_stacktrace=1")
    return subprocess.run([sys.executable, "-m", "pytest", "-x", "-q", "-s", "-p", "no:cacheprovider"] + args, cwd=ROOT, env=env,
                          capture_output=True, text=True, timeout=timeout)


def test_golden_and_probe_vectors_under_asan_ubsan():
    p = sanitized_pytest(["tests/test_emul_parity.py", "-k", "golden_fixtures or probe_vectors",
                          "tests/test_emul_stream.py::test_stream_wasm_pattern"], 900)
    tail = (p.stdout + p.stderr)[-3000:]
    assert p.returncode == 0, tail
    assert "runtime error" not in p.stdout + p.stderr and "AddressSanitizer" not in p.stdout + p.stderr, tail
