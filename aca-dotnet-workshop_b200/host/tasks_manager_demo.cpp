// tasks_manager_demo.cpp — drives the C ABI exactly as the C# GpuTasksManager would:
// builds the hand-checked eleven-task list that the parity tests use (build_kat), runs the unchanged
// cron handler against GpuTasksManager in both modes and prints what it observed.
// Usage: tasks_manager_demo <path-to-libtasktable.so>
#include <cstdio>

#include "gpu_tasks_manager.hpp"

using namespace taskstracker;

static DateTime dt(int64_t ticks, uint32_t kind) { return DateTime{(uint64_t)ticks | ((uint64_t)kind << 62)}; }

static std::vector<TaskModel> kat() {
  const int64_t today = TT_DEFAULT_TODAY_TICKS, day = TT_TICKS_PER_DAY, yday = today - day;
  struct Row { int64_t due; uint32_t kind; bool completed, overdue; int created; };
  const Row rows[] = {
      {yday, 0, false, false, 500}, {yday + 1, 0, false, false, 100}, {yday, 1, false, false, 200},
      {yday, 2, false, false, 300}, {yday, 0, true, false, 400},      {yday, 0, false, true, 450},
      {yday - day, 0, false, false, 50}, {today, 0, false, false, 60}, {yday, 0, false, false, 10},
      {yday, 0, false, false, 500}, {yday + 36000000000LL, 1, false, false, 70}};
  std::vector<TaskModel> out;
  int i = 0;
  for (const Row& r : rows) {
    TaskModel t;
    const uint8_t b = (uint8_t)(i + 1);
    t.TaskId = Guid{b, b, b, b, b, b, 0x40, b, 0x80, b, b, b, b, b, b, b};
    t.TaskName = "Task " + std::to_string(i);
    t.TaskCreatedBy = i % 2 ? "bob@x.io" : "ann@x.io";
    t.TaskAssignedTo = "x@y.z";
    t.TaskCreatedOn = dt(today - 30 * day + r.created, 1);
    t.TaskDueDate = dt(r.due, r.kind);
    t.IsCompleted = r.completed;
    t.IsOverDue = r.overdue;
    out.push_back(t);
    ++i;
  }
  return out;
}

int main(int argc, char** argv) {
  if (argc < 2) {
    fprintf(stderr, "usage: %s <libtasktable.so>\n", argv[0]);
    return 2;
  }
  try {
    auto lib = std::make_shared<TaskTableLibrary>(argv[1]);
    Clock clock;
    for (int32_t mode : {TT_MODE_STORE, TT_MODE_FAKE}) {
      GpuTasksManager manager(lib, kat(), mode, &clock);
      printf("%s due:", mode == TT_MODE_STORE ? "STORE" : "FAKE");
      for (const TaskModel& t : manager.GetYesterdaysDueTasks()) printf(" %s", t.TaskName.c_str());
      printf("\n");
      if (mode == TT_MODE_STORE) printf("GET api/overduetasks: %s\n", manager.OverdueTasksGet().c_str());
      try {
        size_t n = CheckOverDueTasksJob(manager, clock);
        printf("cron marked %zu\n", n);
      } catch (const NotImplementedException& e) {
        printf("cron threw NotImplementedException\n");
      }
      printf("after:");
      for (const TaskModel& t : manager.GetYesterdaysDueTasks()) printf(" %s", t.TaskName.c_str());
      printf("\nann:");
      for (const TaskModel& t : manager.GetTasksByCreator("ann@x.io")) printf(" %s%s", t.TaskName.c_str(), t.IsOverDue ? "*" : "");
      printf("\n");
      // the single-row methods of the interface
      Guid id = manager.CreateNewTask("Task new", "ann@x.io", "q@y.z", dt(TT_DEFAULT_TODAY_TICKS - TT_TICKS_PER_DAY, 0));
      TaskModel got;
      bool found = manager.GetTaskById(id, &got);
      bool done = manager.MarkTaskCompleted(id);
      manager.GetTaskById(id, &got);
      printf("new: found=%d name=%s completed=%d", found, got.TaskName.c_str(), got.IsCompleted);
      bool del = manager.DeleteTask(id), again = manager.DeleteTask(id);
      printf(" marked=%d deleted=%d again=%d gone=%d\n", done, del, again, !manager.GetTaskById(id, nullptr));
    }
  } catch (const std::exception& e) {
    fprintf(stderr, "error: %s\n", e.what());
    return 1;
  }
  return 0;
}
