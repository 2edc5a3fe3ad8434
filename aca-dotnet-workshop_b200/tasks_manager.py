"""GpuTasksManager — the reference's ITasksManager, served from the HBM-resident task table.

Host-side mirror of Api/Services/ITasksManager.cs:5-15 for the three methods on the hot path
(GetYesterdaysDueTasks :14, MarkOverdueTasks :13, GetTasksByCreator :7) plus the cron
handler that drives them (Svc/Controllers/ScheduledTasksManagerController.cs:19-46).  Method
names, argument meaning and error behaviour follow the reference: FAKE mode's
MarkOverdueTasks raises (NotImplementedException, Api/Services/FakeTasksManager.cs:102-105);
STORE mode follows Api/Services/TasksStoreManager.cs.  The C# wrapper in INTEGRATION.md has
the same shape over the same C ABI.

Only columns a kernel reads live on the device; TaskName / TaskAssignedTo stay here, keyed
by row, and are re-attached when TaskModel objects are rebuilt from a result.
"""
import uuid
from dataclasses import dataclass
from typing import List, Optional

import numpy as np

from . import tasktable as tt

KIND_UNSPECIFIED, KIND_UTC, KIND_LOCAL = 0, 1, 2
_TICKS_MASK = (1 << 62) - 1


@dataclass
class TaskModel:
    """Api/Models/TaskModel.cs:3-13.  DateTime fields are .NET dateData (ticks | kind << 62)."""

    TaskId: uuid.UUID
    TaskName: str = ""
    TaskCreatedBy: str = ""
    TaskCreatedOn: int = 0
    TaskDueDate: int = 0
    TaskAssignedTo: str = ""
    IsCompleted: bool = False
    IsOverDue: bool = False


def date_data(ticks, kind=KIND_UNSPECIFIED):
    return (ticks & _TICKS_MASK) | (kind << 62)


class GpuTasksManager:
    """Drop-in for FakeTasksManager / TasksStoreManager on the sweep and creator-filter path.

    mode = tt.MODE_FAKE or tt.MODE_STORE selects which reference implementation's semantics
    are reproduced (they differ: SURVEY.md §0).  The clock is injected: `today_local_ticks`
    stands for DateTime.Today, `run_at_utc_ticks` for DateTime.UtcNow in the cron handler.
    """

    def __init__(self, tasks: List[TaskModel], mode=tt.MODE_STORE, device=0, new_guid=uuid.uuid4, utc_now=None):
        self.mode = mode
        self._new_guid = new_guid            # Guid.NewGuid()
        self._utc_now = utc_now or (lambda: date_data(tt.DEFAULT_RUN_AT, KIND_UTC))  # DateTime.UtcNow
        self._names = [t.TaskName for t in tasks]
        self._assignees = [t.TaskAssignedTo for t in tasks]
        creators = {}
        refs = np.empty(len(tasks), dtype=np.uint32)
        for i, t in enumerate(tasks):
            b = t.TaskCreatedBy.encode("utf-8")
            refs[i] = creators.setdefault(b, len(creators))
        self._creators = list(creators.keys()) or [b""]  # mirrors the library's dictionary, index for index
        ids = np.frombuffer(b"".join(t.TaskId.bytes for t in tasks), dtype=np.uint8).reshape(-1, 16)
        self._table = tt.TaskTable(device).load_soa(
            ids,
            np.array([t.TaskCreatedOn for t in tasks], dtype=np.uint64),
            np.array([t.TaskDueDate for t in tasks], dtype=np.uint64),
            np.array([t.IsCompleted for t in tasks], dtype=np.uint8),
            np.array([t.IsOverDue for t in tasks], dtype=np.uint8),
            refs,
            self._creators,
        )
        if tasks:
            self._table.load_payload([n.encode("utf-8") for n in self._names], [a.encode("utf-8") for a in self._assignees])

    @property
    def table(self):
        return self._table

    def _materialise(self, result) -> List[TaskModel]:
        rows = result.rows()
        ids = result.ids()
        created = result.field(tt.FIELD_CREATED_RAW)
        due = result.field(tt.FIELD_DUE_RAW)
        comp = result.field(tt.FIELD_COMPLETED)
        over = result.field(tt.FIELD_OVERDUE)
        ref = result.field(tt.FIELD_CREATOR_REF)
        out = []
        for k, r in enumerate(rows.tolist()):
            out.append(TaskModel(uuid.UUID(bytes=ids[k].tobytes()), self._names[r], self._creator_name(int(ref[k])),
                                 int(created[k]), int(due[k]), self._assignees[r], bool(comp[k]), bool(over[k])))
        result.free()
        return out

    def _creator_name(self, ref):
        if ref == 0xFFFFFFFE:  # the stored document said "taskCreatedBy": null
            return None
        return self._creators[ref].decode("utf-8")

    # ITasksManager.cs:8
    def GetTaskById(self, taskId: uuid.UUID) -> Optional[TaskModel]:
        found = self._materialise(self._table.get_by_id(taskId.bytes))
        return found[0] if found else None

    # ITasksManager.cs:9 — FakeTasksManager.cs:32-47 / TasksStoreManager.cs:22-38
    def CreateNewTask(self, taskName: str, createdBy: str, assignedTo: str, dueDate: int) -> uuid.UUID:
        task_id = self._new_guid()
        b = createdBy.encode("utf-8")
        if b not in self._creator_set():
            self._creators.append(b)  # the library appends new creators in the same order
            self._creator_lookup.add(b)
        row = self._table.insert_task(task_id.bytes, self._utc_now(), dueDate, b)
        assert row == len(self._names)
        self._table.set_task_strings(task_id.bytes, taskName.encode("utf-8"), assignedTo.encode("utf-8"))
        self._names.append(taskName)
        self._assignees.append(assignedTo)
        return task_id

    def _creator_set(self):
        if not hasattr(self, "_creator_lookup"):
            self._creator_lookup = set(self._creators)
        return self._creator_lookup

    # ITasksManager.cs:10 — FakeTasksManager.cs:87-100 / TasksStoreManager.cs:84-102
    def UpdateTask(self, taskId: uuid.UUID, taskName: str, assignedTo: str, dueDate: int) -> bool:
        res = self._table.get_by_id(taskId.bytes)
        rows = res.rows()
        res.free()
        if rows.size == 0:
            if self.mode == tt.MODE_STORE:
                # TasksStoreManager.cs:88 dereferences the missing document before its null check
                raise AttributeError("NullReferenceException: Object reference not set to an instance of an object")
            return False
        self._names[int(rows[0])] = taskName
        self._assignees[int(rows[0])] = assignedTo
        self._table.set_task_strings(taskId.bytes, taskName.encode("utf-8"), assignedTo.encode("utf-8"))
        return self._table.update_task(taskId.bytes, dueDate)

    # ITasksManager.cs:11
    def MarkTaskCompleted(self, taskId: uuid.UUID) -> bool:
        return self._table.mark_completed(taskId.bytes)

    # ITasksManager.cs:12 — the state-store version reports true whether or not the key existed
    def DeleteTask(self, taskId: uuid.UUID) -> bool:
        found = self._table.delete_task(taskId.bytes)
        return True if self.mode == tt.MODE_STORE else found

    # ITasksManager.cs:14
    def GetYesterdaysDueTasks(self, today_local_ticks=tt.DEFAULT_TODAY) -> List[TaskModel]:
        return self._materialise(self._table.get_yesterdays_due(self.mode, today_local_ticks))

    # ITasksManager.cs:13 — TasksStoreManager.cs:141-149: IsOverDue = true, then an upsert of the whole
    # record under its TaskId.  The listed tasks that exist get their bit set in one kernel; an id the
    # table does not hold is created from the caller's record, as SaveStateAsync would.
    def MarkOverdueTasks(self, overDueTasksList: List[TaskModel]) -> None:
        ids = np.frombuffer(b"".join(t.TaskId.bytes for t in overDueTasksList), dtype=np.uint8).reshape(-1, 16)
        if self.mode == tt.MODE_FAKE:
            self._table.mark_overdue(self.mode, ids)  # raises NotImplementedInReference
        for t in overDueTasksList:  # TasksStoreManager.cs:146 mutates the caller's objects too
            t.IsOverDue = True
        if not overDueTasksList:
            return
        _, missing = self._table.mark_overdue(self.mode, ids)
        self.last_mark_missing = missing
        if missing:
            for t in overDueTasksList:
                res = self._table.get_by_id(t.TaskId.bytes)
                known = res.count > 0
                res.free()
                if not known:
                    self._upsert(t)

    def _upsert(self, t: TaskModel) -> None:
        b = t.TaskCreatedBy.encode("utf-8")
        if b not in self._creator_set():
            self._creators.append(b)
            self._creator_lookup.add(b)
        row, existed = self._table.upsert_task(t.TaskId.bytes, t.TaskCreatedOn, t.TaskDueDate, b, t.IsCompleted, t.IsOverDue)
        if not existed:
            assert row == len(self._names)
            self._names.append(t.TaskName)
            self._assignees.append(t.TaskAssignedTo)
        else:
            self._names[row], self._assignees[row] = t.TaskName, t.TaskAssignedTo
        self._table.set_task_strings(t.TaskId.bytes, t.TaskName.encode("utf-8"), t.TaskAssignedTo.encode("utf-8"))

    # ITasksManager.cs:7
    def GetTasksByCreator(self, createdBy: str) -> List[TaskModel]:
        return self._materialise(self._table.get_by_creator(createdBy))

    # What the two list endpoints send, written on the device (tt_result_json): the body of
    # OverdueTasksController.Get (Api/Controllers/OverdueTasksController.cs:22-26) and of
    # TasksController.Get (Api/Controllers/TasksController.cs:21-25); no TaskModel objects are built.
    def OverdueTasksGet(self, today_local_ticks=tt.DEFAULT_TODAY, local_offset_min=0) -> bytes:
        return self._json(self._table.get_yesterdays_due(self.mode, today_local_ticks), local_offset_min)

    def TasksGet(self, createdBy: str, local_offset_min=0) -> bytes:
        return self._json(self._table.get_by_creator(createdBy), local_offset_min)

    @staticmethod
    def _json(result, local_offset_min):
        try:
            return result.to_json(tt.JSON_RESPONSE, local_offset_min)
        finally:
            result.free()

    # Svc/Controllers/ScheduledTasksManagerController.cs:19-46, unfused: three calls like the reference
    def CheckOverDueTasksJob(self, run_at_utc_ticks=tt.DEFAULT_RUN_AT, today_local_ticks=tt.DEFAULT_TODAY) -> int:
        tasksList = self.GetYesterdaysDueTasks(today_local_ticks)
        run_date = (run_at_utc_ticks & _TICKS_MASK) // tt.TICKS_PER_DAY
        overdueTasksList = [t for t in tasksList if run_date > (t.TaskDueDate & _TICKS_MASK) // tt.TICKS_PER_DAY]
        if len(overdueTasksList) > 0:
            self.MarkOverdueTasks(overdueTasksList)
        return len(overdueTasksList)

    # the same tick as one device pass (tt_sweep)
    def SweepOverdueTasks(self, run_at_utc_ticks=tt.DEFAULT_RUN_AT, today_local_ticks=tt.DEFAULT_TODAY):
        result, n_marked, threw = self._table.sweep(self.mode, today_local_ticks, run_at_utc_ticks)
        tasks = self._materialise(result)
        if threw:
            raise tt.NotImplementedInReference(tt.TT_ERR_NOT_IMPLEMENTED, "MarkOverdueTasks (FAKE mode)")
        return tasks, n_marked

    def close(self):
        self._table.close()
