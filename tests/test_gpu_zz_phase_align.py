"""Scheduling knob of the team kernel (SRT_TEAM_PHASE_ALIGN, srt_team.cuh): results must not depend on it.
Named to run last and executed in a child process with a time limit: the knob changes when idle teams take rays from the
queue, so a mistake would show as a hang, not as a wrong number."""
import os
import subprocess
import sys

import pytest

import common

pytestmark = pytest.mark.gpu


def test_phase_align_is_pure_scheduling(built):
    script = os.path.join(common.ROOT, "tests", "phase_align_check.py")
    r = subprocess.run([sys.executable, script], capture_output=True, text=True, timeout=240, cwd=common.ROOT)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert "PHASE_ALIGN_OK" in r.stdout
