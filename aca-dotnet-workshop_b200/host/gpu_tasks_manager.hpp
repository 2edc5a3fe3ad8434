// gpu_tasks_manager.hpp — C++ host mirror of the reference's ITasksManager over the C ABI.
//
// The reference's host language is C# (no .NET toolchain exists on this box), so the host side
// above the C ABI is written in C++ with the reference's own names and error behaviour:
//   ITasksManager                Api/Services/ITasksManager.cs:5-15
//   TaskModel                    Api/Models/TaskModel.cs:3-13
//   GetYesterdaysDueTasks        FakeTasksManager.cs:107-112 / TasksStoreManager.cs:104-139
//   MarkOverdueTasks             FakeTasksManager.cs:102-105 (throws) / TasksStoreManager.cs:141-149
//   GetTasksByCreator            FakeTasksManager.cs:68-72 / TasksStoreManager.cs:54-69
//   CheckOverDueTasksJob         Svc/Controllers/ScheduledTasksManagerController.cs:19-46
// The C# version of this class binds the same entry points with [DllImport].
// libtasktable.so is loaded with dlopen so this header builds with a plain C++ compiler.
#pragma once

#include <dlfcn.h>

#include <array>
#include <cstdint>
#include <cstring>
#include <map>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>

#include "../../include/tasktable.h"

namespace taskstracker {

using Guid = std::array<uint8_t, 16>;  // canonical big-endian bytes of Guid.ToString("D")

struct DateTime {  // .NET dateData: ticks | kind << 62
  uint64_t dateData = 0;
  int64_t Ticks() const { return (int64_t)(dateData & TT_TICKS_MASK); }
  uint32_t Kind() const { return (uint32_t)(dateData >> 62); }
  DateTime Date() const { int64_t t = Ticks(); return DateTime{(uint64_t)(t - t % TT_TICKS_PER_DAY) | ((uint64_t)Kind() << 62)}; }
  bool operator>(const DateTime& o) const { return Ticks() > o.Ticks(); }
};

struct TaskModel {
  Guid TaskId{};
  std::string TaskName;
  std::string TaskCreatedBy;
  DateTime TaskCreatedOn;
  DateTime TaskDueDate;
  std::string TaskAssignedTo;
  bool IsCompleted = false;
  bool IsOverDue = false;
};

struct NotImplementedException : std::logic_error {
  NotImplementedException() : std::logic_error("The method or operation is not implemented.") {}
};

struct TasksManagerException : std::runtime_error {
  int32_t code;
  TasksManagerException(int32_t c, const std::string& m) : std::runtime_error(m), code(c) {}
};

// Api/Services/ITasksManager.cs:5-15, method for method (Task<T> becomes T: the calls block).
struct ITasksManager {
  virtual ~ITasksManager() = default;
  virtual std::vector<TaskModel> GetTasksByCreator(const std::string& createdBy) = 0;
  virtual bool GetTaskById(const Guid& taskId, TaskModel* out) = 0;  // false = the reference's null
  virtual Guid CreateNewTask(const std::string& taskName, const std::string& createdBy, const std::string& assignedTo,
                             DateTime dueDate) = 0;
  virtual bool UpdateTask(const Guid& taskId, const std::string& taskName, const std::string& assignedTo,
                          DateTime dueDate) = 0;
  virtual bool MarkTaskCompleted(const Guid& taskId) = 0;
  virtual bool DeleteTask(const Guid& taskId) = 0;
  virtual void MarkOverdueTasks(std::vector<TaskModel>& overDueTasksList) = 0;
  virtual std::vector<TaskModel> GetYesterdaysDueTasks() = 0;
};

// Injected clock: the reference reads DateTime.Today / DateTime.UtcNow from the environment.
struct Clock {
  int64_t today_local_ticks = TT_DEFAULT_TODAY_TICKS;
  int64_t utc_now_ticks = TT_DEFAULT_RUN_AT_TICKS;
};

class TaskTableLibrary {
 public:
  explicit TaskTableLibrary(const std::string& path) {
    h_ = dlopen(path.c_str(), RTLD_NOW | RTLD_LOCAL);
    if (!h_) throw TasksManagerException(TT_ERR_STATE, std::string("cannot load ") + path + ": " + dlerror());
#define TT_BIND(name) name = reinterpret_cast<decltype(&::name)>(must(#name))
    TT_BIND(tt_create); TT_BIND(tt_destroy); TT_BIND(tt_last_error); TT_BIND(tt_load_soa);
    TT_BIND(tt_get_yesterdays_due); TT_BIND(tt_mark_overdue); TT_BIND(tt_sweep); TT_BIND(tt_get_by_creator);
    TT_BIND(tt_result_count); TT_BIND(tt_result_rows); TT_BIND(tt_result_ids); TT_BIND(tt_result_field);
    TT_BIND(tt_result_free);
    TT_BIND(tt_get_by_id); TT_BIND(tt_insert_task); TT_BIND(tt_update_task); TT_BIND(tt_mark_completed);
    TT_BIND(tt_delete_task); TT_BIND(tt_upsert_task);
    TT_BIND(tt_load_payload); TT_BIND(tt_set_task_strings); TT_BIND(tt_result_json); TT_BIND(tt_json_size); TT_BIND(tt_json_copy); TT_BIND(tt_json_free);
#undef TT_BIND
  }
  ~TaskTableLibrary() { if (h_) dlclose(h_); }
  decltype(&::tt_create) tt_create;
  decltype(&::tt_destroy) tt_destroy;
  decltype(&::tt_last_error) tt_last_error;
  decltype(&::tt_load_soa) tt_load_soa;
  decltype(&::tt_get_yesterdays_due) tt_get_yesterdays_due;
  decltype(&::tt_mark_overdue) tt_mark_overdue;
  decltype(&::tt_sweep) tt_sweep;
  decltype(&::tt_get_by_creator) tt_get_by_creator;
  decltype(&::tt_result_count) tt_result_count;
  decltype(&::tt_result_rows) tt_result_rows;
  decltype(&::tt_result_ids) tt_result_ids;
  decltype(&::tt_result_field) tt_result_field;
  decltype(&::tt_result_free) tt_result_free;
  decltype(&::tt_get_by_id) tt_get_by_id;
  decltype(&::tt_insert_task) tt_insert_task;
  decltype(&::tt_update_task) tt_update_task;
  decltype(&::tt_mark_completed) tt_mark_completed;
  decltype(&::tt_delete_task) tt_delete_task;
  decltype(&::tt_upsert_task) tt_upsert_task;
  decltype(&::tt_load_payload) tt_load_payload;
  decltype(&::tt_set_task_strings) tt_set_task_strings;
  decltype(&::tt_result_json) tt_result_json;
  decltype(&::tt_json_size) tt_json_size;
  decltype(&::tt_json_copy) tt_json_copy;
  decltype(&::tt_json_free) tt_json_free;

 private:
  void* must(const char* name) {
    void* p = dlsym(h_, name);
    if (!p) throw TasksManagerException(TT_ERR_STATE, std::string("missing symbol ") + name);
    return p;
  }
  void* h_ = nullptr;
};

class GpuTasksManager : public ITasksManager {
 public:
  // mode: TT_MODE_FAKE reproduces FakeTasksManager, TT_MODE_STORE reproduces TasksStoreManager
  GpuTasksManager(std::shared_ptr<TaskTableLibrary> lib, const std::vector<TaskModel>& tasks, int32_t mode,
                  Clock* clock, int32_t device = 0)
      : lib_(std::move(lib)), mode_(mode), clock_(clock) {
    const int64_t n = (int64_t)tasks.size();
    std::vector<uint8_t> ids(16 * n), comp(n), over(n);
    std::vector<uint64_t> created(n), due(n);
    std::vector<uint32_t> ref(n), off{0};
    std::vector<uint8_t> pool;
    std::map<std::string, uint32_t> dict;
    for (int64_t i = 0; i < n; ++i) {
      const TaskModel& t = tasks[i];
      memcpy(&ids[16 * i], t.TaskId.data(), 16);
      created[i] = t.TaskCreatedOn.dateData;
      due[i] = t.TaskDueDate.dateData;
      comp[i] = t.IsCompleted;
      over[i] = t.IsOverDue;
      auto it = dict.find(t.TaskCreatedBy);
      if (it == dict.end()) {
        it = dict.emplace(t.TaskCreatedBy, (uint32_t)creators_.size()).first;
        creators_.push_back(t.TaskCreatedBy);
        pool.insert(pool.end(), t.TaskCreatedBy.begin(), t.TaskCreatedBy.end());
        off.push_back((uint32_t)pool.size());
      }
      ref[i] = it->second;
      names_.push_back(t.TaskName);
      assignees_.push_back(t.TaskAssignedTo);
    }
    if (creators_.empty()) { creators_.push_back(""); off.push_back(0); }
    if (pool.empty()) pool.push_back(0);
    check(lib_->tt_create(device, &table_));
    tt_soa_view v{n, ids.data(), created.data(), due.data(), comp.data(), over.data(), ref.data(),
                  (uint32_t)creators_.size(), off.data(), pool.data()};
    check(lib_->tt_load_soa(table_, &v));
    // TaskName / TaskAssignedTo go to the device too (dictionary-encoded), so that a controller can send
    // the device-made response text instead of materialising TaskModel objects (OverdueTasksGet below)
    Encoded names = encode(names_), assignees = encode(assignees_);
    tt_strings_view nv{names.ref.data(), (uint32_t)(names.off.size() - 1), names.off.data(), names.pool.data()};
    tt_strings_view av{assignees.ref.data(), (uint32_t)(assignees.off.size() - 1), assignees.off.data(), assignees.pool.data()};
    check(lib_->tt_load_payload(table_, &nv, &av));
  }
  ~GpuTasksManager() override { if (table_) lib_->tt_destroy(table_); }

  std::vector<TaskModel> GetYesterdaysDueTasks() override {
    tt_result* r = nullptr;
    check(lib_->tt_get_yesterdays_due(table_, mode_, clock_->today_local_ticks, &r));
    return materialise(r);
  }

  void MarkOverdueTasks(std::vector<TaskModel>& overDueTasksList) override {
    if (mode_ == TT_MODE_FAKE) throw NotImplementedException();  // FakeTasksManager.cs:104
    std::vector<uint8_t> ids(16 * overDueTasksList.size());
    for (size_t i = 0; i < overDueTasksList.size(); ++i) {
      overDueTasksList[i].IsOverDue = true;  // TasksStoreManager.cs:146
      memcpy(&ids[16 * i], overDueTasksList[i].TaskId.data(), 16);
    }
    int64_t marked = 0, missing = 0;
    check(lib_->tt_mark_overdue(table_, mode_, ids.data(), (int64_t)overDueTasksList.size(), &marked, &missing));
    last_mark_missing = missing;
    if (missing == 0) return;
    // SaveStateAsync is an upsert (TasksStoreManager.cs:147): a listed task the table does not hold is created
    for (const TaskModel& t : overDueTasksList) {
      tt_result* r = nullptr;
      check(lib_->tt_get_by_id(table_, t.TaskId.data(), &r));
      int64_t n = 0;
      lib_->tt_result_count(r, &n);
      lib_->tt_result_free(r);
      if (n > 0) continue;
      int64_t row = -1;
      int32_t existed = 0;
      check(lib_->tt_upsert_task(table_, t.TaskId.data(), t.TaskCreatedOn.dateData, t.TaskDueDate.dateData,
                                 t.TaskCreatedBy.data(), (int32_t)t.TaskCreatedBy.size(), t.IsCompleted ? 1 : 0, 1, &row,
                                 &existed));
      int32_t found = 0;
      check(lib_->tt_set_task_strings(table_, t.TaskId.data(), t.TaskName.data(), (int32_t)t.TaskName.size(),
                                      t.TaskAssignedTo.data(), (int32_t)t.TaskAssignedTo.size(), &found));
    }
  }
  int64_t last_mark_missing = 0;  // ids of the last MarkOverdueTasks call that had to be created

  std::vector<TaskModel> GetTasksByCreator(const std::string& createdBy) override {
    tt_result* r = nullptr;
    check(lib_->tt_get_by_creator(table_, createdBy.data(), (int32_t)createdBy.size(), &r));
    return materialise(r);
  }

  // ---- the single-row methods (FakeTasksManager.cs:32-100 / TasksStoreManager.cs:22-102) ----
  bool GetTaskById(const Guid& taskId, TaskModel* out) override {
    tt_result* r = nullptr;
    check(lib_->tt_get_by_id(table_, taskId.data(), &r));
    std::vector<TaskModel> found = materialise(r);
    if (found.empty()) return false;
    if (out) *out = found[0];
    return true;
  }

  Guid CreateNewTask(const std::string& taskName, const std::string& createdBy, const std::string& assignedTo,
                     DateTime dueDate) override {
    Guid id = NewGuid();
    int64_t row = -1;
    check(lib_->tt_insert_task(table_, id.data(), (uint64_t)clock_->utc_now_ticks | (1ull << 62), dueDate.dateData,
                               createdBy.data(), (int32_t)createdBy.size(), &row));
    bool known = false;
    for (const std::string& c : creators_) known = known || c == createdBy;
    if (!known) creators_.push_back(createdBy);  // the library appends new creators in the same order
    names_.push_back(taskName);
    assignees_.push_back(assignedTo);
    check(lib_->tt_set_task_strings(table_, id.data(), taskName.data(), (int32_t)taskName.size(), assignedTo.data(),
                                    (int32_t)assignedTo.size(), nullptr));
    return id;
  }

  bool UpdateTask(const Guid& taskId, const std::string& taskName, const std::string& assignedTo,
                  DateTime dueDate) override {
    tt_result* r = nullptr;
    check(lib_->tt_get_by_id(table_, taskId.data(), &r));
    int64_t n = 0;
    uint32_t row = 0;
    check(lib_->tt_result_count(r, &n));
    if (n) check(lib_->tt_result_rows(r, &row));
    lib_->tt_result_free(r);
    if (!n) {
      if (mode_ == TT_MODE_STORE)  // TasksStoreManager.cs:88 dereferences the missing document
        throw std::logic_error("NullReferenceException");
      return false;
    }
    names_[row] = taskName;
    assignees_[row] = assignedTo;
    check(lib_->tt_set_task_strings(table_, taskId.data(), taskName.data(), (int32_t)taskName.size(), assignedTo.data(),
                                    (int32_t)assignedTo.size(), nullptr));
    int32_t found = 0;
    check(lib_->tt_update_task(table_, taskId.data(), dueDate.dateData, &found));
    return found != 0;
  }

  bool MarkTaskCompleted(const Guid& taskId) override {
    int32_t found = 0;
    check(lib_->tt_mark_completed(table_, taskId.data(), &found));
    return found != 0;
  }

  bool DeleteTask(const Guid& taskId) override {
    int32_t found = 0;
    check(lib_->tt_delete_task(table_, taskId.data(), &found));
    return mode_ == TT_MODE_STORE ? true : found != 0;  // DeleteStateAsync answers true either way
  }

  // The whole tick as one device pass (tt_sweep).  Returns the listed tasks; *n_marked = rows set.
  // The body OverdueTasksController.Get (Api/Controllers/OverdueTasksController.cs:22-26) sends for
  // `return await _tasksManager.GetYesterdaysDueTasks();`, written on the device: no TaskModel objects
  // are built on the host (CreateNewTask / UpdateTask keep the device's strings current).
  std::string OverdueTasksGet(int32_t local_offset_min = 0) {
    tt_result* r = nullptr;
    check(lib_->tt_get_yesterdays_due(table_, mode_, clock_->today_local_ticks, &r));
    tt_json* j = nullptr;
    int32_t rc = lib_->tt_result_json(r, TT_JSON_RESPONSE, local_offset_min, &j);
    lib_->tt_result_free(r);
    check(rc);
    int64_t n = 0;
    check(lib_->tt_json_size(j, &n));
    std::string body((size_t)n, '\0');
    rc = lib_->tt_json_copy(j, &body[0]);
    lib_->tt_json_free(j);
    check(rc);
    return body;
  }

  std::vector<TaskModel> SweepOverdueTasks(int64_t* n_marked) {
    tt_result* r = nullptr;
    int32_t rc = lib_->tt_sweep(table_, mode_, clock_->today_local_ticks, clock_->utc_now_ticks, &r, n_marked);
    if (rc != TT_OK && rc != TT_ERR_NOT_IMPLEMENTED) check(rc);
    std::vector<TaskModel> out = materialise(r);
    if (rc == TT_ERR_NOT_IMPLEMENTED) throw NotImplementedException();
    return out;
  }

 private:
  // Guid.NewGuid(): RFC 4122 version 4 from a splitmix stream (deterministic for the demo)
  Guid NewGuid() {
    Guid g;
    for (int half = 0; half < 2; ++half) {
      uint64_t v = tt_mix64(guid_state_++);
      for (int i = 0; i < 8; ++i) g[8 * half + i] = (uint8_t)(v >> (8 * i));
    }
    g[6] = (uint8_t)((g[6] & 0x0F) | 0x40);
    g[8] = (uint8_t)((g[8] & 0x3F) | 0x80);
    return g;
  }
  uint64_t guid_state_ = 0x1234;

  struct Encoded {
    std::vector<uint32_t> ref, off{0};
    std::vector<uint8_t> pool;
  };
  static Encoded encode(const std::vector<std::string>& values) {
    Encoded e;
    std::map<std::string, uint32_t> dict;
    for (const std::string& s : values) {
      auto it = dict.find(s);
      if (it == dict.end()) {
        it = dict.emplace(s, (uint32_t)(e.off.size() - 1)).first;
        e.pool.insert(e.pool.end(), s.begin(), s.end());
        e.off.push_back((uint32_t)e.pool.size());
      }
      e.ref.push_back(it->second);
    }
    if (e.pool.empty()) e.pool.push_back(0);
    return e;
  }

  void check(int32_t rc) {
    if (rc == TT_OK) return;
    if (rc == TT_ERR_NOT_IMPLEMENTED) throw NotImplementedException();
    throw TasksManagerException(rc, lib_->tt_last_error());
  }

  std::vector<TaskModel> materialise(tt_result* r) {
    int64_t n = 0;
    check(lib_->tt_result_count(r, &n));
    std::vector<uint32_t> rows(n), ref(n);
    std::vector<uint8_t> ids(16 * n), comp(n), over(n);
    std::vector<uint64_t> created(n), due(n);
    check(lib_->tt_result_rows(r, rows.data()));
    check(lib_->tt_result_ids(r, ids.data()));
    check(lib_->tt_result_field(r, TT_FIELD_CREATED_RAW, created.data()));
    check(lib_->tt_result_field(r, TT_FIELD_DUE_RAW, due.data()));
    check(lib_->tt_result_field(r, TT_FIELD_COMPLETED, comp.data()));
    check(lib_->tt_result_field(r, TT_FIELD_OVERDUE, over.data()));
    check(lib_->tt_result_field(r, TT_FIELD_CREATOR_REF, ref.data()));
    lib_->tt_result_free(r);
    std::vector<TaskModel> out(n);
    for (int64_t i = 0; i < n; ++i) {
      memcpy(out[i].TaskId.data(), &ids[16 * i], 16);
      out[i].TaskName = names_[rows[i]];
      out[i].TaskCreatedBy = ref[i] < creators_.size() ? creators_[ref[i]] : std::string();  // 0xFFFFFFFE: a null string
      out[i].TaskCreatedOn.dateData = created[i];
      out[i].TaskDueDate.dateData = due[i];
      out[i].TaskAssignedTo = assignees_[rows[i]];
      out[i].IsCompleted = comp[i] != 0;
      out[i].IsOverDue = over[i] != 0;
    }
    return out;
  }

  std::shared_ptr<TaskTableLibrary> lib_;
  int32_t mode_;
  Clock* clock_;
  tt_table* table_ = nullptr;
  std::vector<std::string> names_, assignees_, creators_;
};

// Svc/Controllers/ScheduledTasksManagerController.cs:19-46 — the cron handler, unchanged in
// shape, calling whatever ITasksManager it is given.  Returns the number of tasks it marked.
inline size_t CheckOverDueTasksJob(ITasksManager& manager, const Clock& clock) {
  DateTime runAt{(uint64_t)clock.utc_now_ticks | (1ull << 62)};                    // :22 DateTime.UtcNow
  std::vector<TaskModel> overdueTasksList;
  std::vector<TaskModel> tasksList = manager.GetYesterdaysDueTasks();              // :28
  for (const TaskModel& taskModel : tasksList)                                     // :32-38
    if (runAt.Date() > taskModel.TaskDueDate.Date()) overdueTasksList.push_back(taskModel);
  if (overdueTasksList.size() > 0) manager.MarkOverdueTasks(overdueTasksList);     // :40-45
  return overdueTasksList.size();
}

}  // namespace taskstracker
