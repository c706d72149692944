"""Build-level guard against the stall of DESIGN.md 3.4: in one code generation the two loop-top ballots of k_wf_tail were
compiled, at the loop's back edge, to bare VOTE instructions (no BRA.DIV divergence check, no WARPSYNC in front), and the
kernel deadlocked on the B200. The SASS of the freshly built objects is scanned with scripts/sass_collectives.py: the
persistent-loop kernels must not contain a ballot without such a guard. Runs without a GPU (cuobjdump reads the object)."""
import os
import shutil
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OBJ = os.path.join(ROOT, "raytracer-cuda_b200", "build", "crt_wavefront.o")


@pytest.mark.skipif(shutil.which("cuobjdump") is None, reason="cuobjdump not on PATH")
def test_loop_ballots_of_the_persistent_kernels_are_guarded():
    if not os.path.exists(OBJ):
        pytest.skip("raytracer-cuda_b200/build/crt_wavefront.o not built (run __graft_entry__.build())")
    out = subprocess.run([sys.executable, os.path.join(ROOT, "scripts", "sass_collectives.py"), OBJ],
                         stdout=subprocess.PIPE, text=True, check=True).stdout
    bare = [l for l in out.splitlines() if ("k_wf_tail" in l or "k_wf_extend" in l) and " VOTE." in l]
    assert not bare, "ballots without BRA.DIV / WARPSYNC in front of them:\n" + "\n".join(bare)
    # the scan does see the kernels (an empty listing because of a renamed symbol would prove nothing)
    sass = subprocess.run(["cuobjdump", "-sass", OBJ], stdout=subprocess.PIPE, text=True).stdout
    assert "k_wf_tail" in sass and "VOTE" in sass
