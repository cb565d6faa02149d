"""GPU: the reference's OWN hot-path test files (tests/indexing/test_hnsw.py, test_ivf.py, tests/inference/test_pq.py,
test_mmap_vector_store.py, tests/services/test_indexing_service.py, test_storage_service.py -- SURVEY.md section 4) run
against the drop-in through the import shim tests/ref_suite_shim.py.  The files live under oracle/_ref/ref_suite/
(git-ignored, fetched by oracle/fetch_ref_suite.py in the build container); without them the test is skipped."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SUITE = os.path.join(ROOT, "oracle", "_ref", "ref_suite", "tests")


@pytest.mark.parametrize("mode", ["full", "hotpath"])
def test_reference_hot_path_suite_passes_against_the_drop_in(mode):
    if not os.path.isdir(SUITE):
        pytest.skip("oracle/_ref/ref_suite not fetched (run oracle/fetch_ref_suite.py where /root/reference exists)")
    env = dict(os.environ, VDB_REF_SHIM_MODE=mode)
    r = subprocess.run([sys.executable, "-m", "pytest", SUITE, "-q", "-p", "no:cacheprovider", "-o", "addopts=", "-x",
                        "--rootdir", SUITE], capture_output=True, text=True, env=env, timeout=1500, cwd=SUITE)
    out_dir = os.path.join(ROOT, "gpurun_out")
    if os.path.isdir(out_dir):
        with open(os.path.join(out_dir, f"ref_suite_{mode}.log"), "w") as f:
            f.write(r.stdout[-20000:] + "\n" + r.stderr[-5000:])
    assert r.returncode == 0, r.stdout[-4000:] + r.stderr[-2000:]
    assert " passed" in r.stdout and " failed" not in r.stdout
