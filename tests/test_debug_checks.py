"""The checked build of the library (-DAKB_DEBUG_CHECKS -DAKB_DEBUG_TRAVERSAL: device-side bounds checks on every node,
leaf-triangle, queue and slot index, trap on traversal-stack overflow) runs the small fixed workload of
tools/sanitize_target.py clean - wavefront with two pipelines, graphs and the finish kernel, megakernel, material-sorted
queues, AO, both BVH builders, ray batches. It stands in for compute-sanitizer, which the GPU pool does not allow."""
import os
import subprocess
import sys

import pytest

from common import ROOT

pytestmark = pytest.mark.gpu


def test_checked_build_runs_the_fixed_workload_clean():
    lib = os.path.join(ROOT, "akarirender-1_b200", "libakari_b200_debug.so")
    assert os.path.exists(lib), "run __graft_entry__.build() first"
    env = dict(os.environ, AKB_LIB=lib)
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "sanitize_target.py")], env=env, capture_output=True,
                       text=True, timeout=600)
    assert r.returncode == 0 and "sanitize target: ok" in r.stdout, r.stdout[-2000:] + r.stderr[-2000:]
