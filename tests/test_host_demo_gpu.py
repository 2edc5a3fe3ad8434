"""The C++ host mirror (GpuTasksManager + the unchanged cron handler) driving the C ABI the way the
C# wrapper would: run the demo binary on the GPU and compare with the hand-derived answers."""
import os
import subprocess

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "aca-dotnet-workshop_b200")

# the GET body is what the oracle's restatement of the MVC serializer prints for Task 8, Task 0, Task 9
EXPECTED = """STORE due: Task 8 Task 0 Task 9
GET api/overduetasks: [{"taskId":"09090909-0909-4009-8009-090909090909","taskName":"Task 8","taskCreatedBy":"ann@x.io","taskCreatedOn":"2026-10-07T00:00:00.000001Z","taskDueDate":"2026-11-05T00:00:00","taskAssignedTo":"x@y.z","isCompleted":false,"isOverDue":false},{"taskId":"01010101-0101-4001-8001-010101010101","taskName":"Task 0","taskCreatedBy":"ann@x.io","taskCreatedOn":"2026-10-07T00:00:00.00005Z","taskDueDate":"2026-11-05T00:00:00","taskAssignedTo":"x@y.z","isCompleted":false,"isOverDue":false},{"taskId":"0a0a0a0a-0a0a-400a-800a-0a0a0a0a0a0a","taskName":"Task 9","taskCreatedBy":"bob@x.io","taskCreatedOn":"2026-10-07T00:00:00.00005Z","taskDueDate":"2026-11-05T00:00:00","taskAssignedTo":"x@y.z","isCompleted":false,"isOverDue":false}]
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
