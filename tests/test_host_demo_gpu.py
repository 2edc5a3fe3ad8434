"""The C++ host mirror (GpuTasksManager + the unchanged cron handler) driving the C ABI the way the
C# wrapper would: run the demo binary on the GPU and compare with the hand-derived answers."""
import os
import subprocess

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "aca-dotnet-workshop_b200")

EXPECTED = """STORE due: Task 8 Task 0 Task 9
cron marked 3
after:
ann: Task 0* Task 4 Task 2 Task 10 Task 6 Task 8*
new: found=1 name=Task new completed=1 marked=1 deleted=1 again=1 gone=1
FAKE due: Task 0 Task 2 Task 3 Task 4 Task 5 Task 8 Task 9
cron threw NotImplementedException
after: Task 0 Task 2 Task 3 Task 4 Task 5 Task 8 Task 9
ann: Task 0 Task 4 Task 2 Task 10 Task 6 Task 8
new: found=1 name=Task new completed=1 marked=1 deleted=1 again=0 gone=1
"""


def test_cpp_manager_and_cron_handler():
    exe = os.path.join(PKG, "host", "tasks_manager_demo")
    if not os.path.exists(exe):
        subprocess.check_call(["make", "-C", ROOT, "aca-dotnet-workshop_b200/host/tasks_manager_demo"])
    out = subprocess.run([exe, os.path.join(PKG, "libtasktable.so")], capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stderr
    assert out.stdout == EXPECTED
