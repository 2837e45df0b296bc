"""The Python side of the `first_gpu_run` GPU tests, run on the host (tools/host_dry_run_gpu_tests.py): the package's Python
API and a host build of the product's own torch binding over the host-emulated C ABI.  By default a subset that finishes
in about two minutes (the densification fixtures through gsr_densify and the binding, the unaligned / masked Adam step,
__graft_entry__.smoke());
GSR_SLOW=1 runs all 25 invocations (about fifteen minutes).  Keeps typos in GPU-only tests from costing GPU minutes."""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
TOOL = os.path.join(ROOT, "tools", "host_dry_run_gpu_tests.py")


def _run(only, expect):
    r = subprocess.run([sys.executable, TOOL, "5000", only], capture_output=True, text=True, timeout=3000)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-4000:]
    assert "dry run complete: %d test invocations" % expect in r.stdout, r.stdout[-2000:]


def test_densification_gpu_tests_python_side_runs_on_the_host():
    _run("reference_models_fixture", 3)          # the three fixtures of the reference's own GaussianModel


def test_adam_gpu_test_with_unaligned_views_and_mask_runs_on_the_host():
    _run("unaligned", 1)


@pytest.mark.skipif(os.environ.get("GSR_SLOW") != "1", reason="all 25 invocations take about fifteen minutes: set GSR_SLOW=1 (green when written)")
def test_every_first_gpu_run_test_covered_by_the_dry_run():
    _run("", 25)


def test_graft_entry_smoke_runs_on_the_host_and_matches_the_oracle():
    """__graft_entry__.smoke() -- the driver's round-end check on cuda:0 -- through the host binding: same Python, same
    binding source, same kernels' logic; sorted lists, ranges, radii, image and gradients against the oracle."""
    # GSR_TILE_SORT=radix: the counting tile sort hands a token from warp to warp, which is one OS-thread wake-up chain per
    # turn in the emulator (2 minutes for this scene); tests/test_abi_emulated.py runs it on the reference build's fixtures
    r = subprocess.run([sys.executable, TOOL, "--scripts", "--smoke"], capture_output=True, text=True, timeout=1500,
                       env=dict(os.environ, GSR_TILE_SORT="radix"))
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-4000:]
    assert "smoke ok" in r.stdout and "ok  __graft_entry__.smoke()" in r.stdout
