"""Build-level guard for the programmatic launches of the single-frame shape ("Latency floor").

A kernel launched programmatically may run before its predecessor has finished; it must not read what the predecessor
writes before `griddepcontrol.wait` (SASS: ACQBULK).  Loads through `const T *__restrict__` kernel parameters are
invariant loads, which nvcc and ptxas both move above the wait -- that happened once (k_orient_describe read the survivor
counts at its first instruction).  The fix routes those reads through pointers laundered after the wait
(borb_cta.cuh); this test looks at the shipped machine code: in the kernels that consume a predecessor's output
through global loads, no global load may precede the first ACQBULK."""
import re
import shutil
import subprocess
from pathlib import Path

import pytest

LIB = Path(__file__).resolve().parents[1] / "orb_slam2_test_b200" / "liborb_b200.so"
# (mangled-name fragment, loads that may stand before the wait)
KERNELS = [("k_quadtreeILi1024", 0),      # reads K2's candidates and counts
           ("k_orient_describeILb1", 0),  # reads K3's survivors and counts (its pattern words are constant data, but
                                          # ptxas places them after the wait anyway)
           ("k_results_to_host", 0)]      # reads the result slab


def _sass(fragment):
    out = subprocess.run(["cuobjdump", "-sass", str(LIB)], capture_output=True, text=True, check=True).stdout
    keep, lines = False, []
    for line in out.splitlines():
        if "Function :" in line:
            keep = fragment in line
        elif keep and re.match(r"\s+/\*[0-9a-f]{4,5}\*/", line):
            lines.append(line)
    return lines


@pytest.mark.skipif(shutil.which("cuobjdump") is None or not LIB.exists(), reason="needs cuobjdump and the built library")
@pytest.mark.parametrize("fragment,allowed", KERNELS)
def test_no_global_load_before_the_grid_dependency_wait(fragment, allowed):
    lines = _sass(fragment)
    assert lines, f"{fragment}: kernel not found in {LIB.name}"
    waits = [i for i, l in enumerate(lines) if "ACQBULK" in l]
    assert waits, f"{fragment}: no griddepcontrol.wait (ACQBULK) in the kernel"
    early = [l.strip() for l in lines[:waits[0]] if re.search(r"\bLDG\b|\bLD\.E\b|\bLDG\.", l)]
    assert len(early) <= allowed, f"{fragment}: global loads before the wait:\n" + "\n".join(early)


@pytest.mark.skipif(shutil.which("cuobjdump") is None or not LIB.exists(), reason="needs cuobjdump and the built library")
def test_the_chain_kernels_release_their_dependents():
    """k_pyr_chain, K2 and K3 let the next kernel be scheduled early (griddepcontrol.launch_dependents: PREEXIT)"""
    for fragment in ("k_pyr_chain", "k_fast_cellsILi64ELi4", "k_quadtreeILi1024"):
        assert any("PREEXIT" in l for l in _sass(fragment)), fragment
