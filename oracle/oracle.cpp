// oracle.cpp — CPU restatement of the reference's overdue-task sweep and creator filter.
//
// TEST INFRASTRUCTURE ONLY.  Nothing in the product path (aca-dotnet-workshop_b200/, include/)
// may include, link, load or call this file; it is the checker used by tests/, by
// __graft_entry__.smoke() and by bench.py's cpu_baseline / --impl reference legs.
//
// PARITY UNPINNED: the reference ships no tests, golden vectors or fixtures for this path
//  and no .NET runtime exists on this box or the GPU box, so this
// restatement cannot be checked against a run of the reference.  It is written line by line
// from the cited C# and from documented .NET 9 BCL behaviour; the only
// pins are the JSON bodies the workshop docs show (docs/aca/05-aca-dapr-pubsubapi/index.md:60-69 and
// three more, tests/golden/reference_doc_samples.json), the seed-row shape
// (Api/Services/FakeTasksManager.cs:12-24), Python's calendar for the date arithmetic and Python's
// JSON reader for the escapes, all covered in tests/test_oracle.py.
//
// Reference files restated here (paths under /root/reference):
//   Api = TasksTracker.TasksManager.Backend.Api, Svc = TasksTracker.Processor.Backend.Svc
//   Api/Models/TaskModel.cs:3-13                     -> struct TaskModel
//   Api/Services/FakeTasksManager.cs:10-25           -> FakeTasksManager::GenerateRandomTasks
//   Api/Services/FakeTasksManager.cs:32-100          -> FakeTasksManager point operations (Create/Delete/GetById/MarkCompleted/Update)
//   Api/Services/TasksStoreManager.cs:22-102         -> TasksStoreManager point operations
//   Api/Services/FakeTasksManager.cs:68-72           -> FakeTasksManager::GetTasksByCreator
//   Api/Services/FakeTasksManager.cs:102-105         -> FakeTasksManager::MarkOverdueTasks (throws)
//   Api/Services/FakeTasksManager.cs:107-112         -> FakeTasksManager::GetYesterdaysDueTasks
//   Api/Services/TasksStoreManager.cs:54-69          -> TasksStoreManager::GetTasksByCreator
//   Api/Services/TasksStoreManager.cs:104-139        -> TasksStoreManager::GetYesterdaysDueTasks
//   Api/Services/TasksStoreManager.cs:141-149        -> TasksStoreManager::MarkOverdueTasks
//   Api/Utilities/DateTimeConverter.cs:26-29         -> FormatConverter
//   Svc/Controllers/ScheduledTasksManagerController.cs:19-46 -> CheckOverDueTasksJob
//
// The data structures are deliberately the reference's: a List<TaskModel> of heap objects
// with UTF-16 strings, scanned sequentially on one thread (LINQ-to-objects is sequential),
// so timing this code is the honest "reference CPU path" baseline.

#include <algorithm>
#include <atomic>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <string>
#include <thread>
#include <unordered_map>
#include <vector>

#include "../include/tt_synth.h"
#include "../include/tt_bmp_defined.h"

namespace {

constexpr int64_t kTicksPerDay = 864000000000LL;
constexpr int64_t kTicksPerSecond = 10000000LL;
constexpr uint64_t kTicksMask = 0x3FFFFFFFFFFFFFFFULL;

// ---- System.DateTime -------------------------------------------
struct DateTime {
  uint64_t dateData = 0;  // ticks | kind << 62
  DateTime() = default;
  explicit DateTime(uint64_t raw) : dateData(raw) {}
  DateTime(int64_t ticks, uint32_t kind) : dateData((uint64_t)ticks | ((uint64_t)kind << 62)) {}
  int64_t Ticks() const { return (int64_t)(dateData & kTicksMask); }
  uint32_t Kind() const { return (uint32_t)(dateData >> 62); }
  // DateTime.Date: time-of-day removed, Kind kept
  DateTime Date() const { int64_t t = Ticks(); return DateTime(t - t % kTicksPerDay, Kind()); }
  DateTime AddDays(int d) const { return DateTime(Ticks() + (int64_t)d * kTicksPerDay, Kind()); }
  DateTime AddMinutes(int m) const { return DateTime(Ticks() + (int64_t)m * 60 * kTicksPerSecond, Kind()); }
  // Equals / operator> / CompareTo compare ticks only and ignore Kind
  bool Equals(const DateTime& o) const { return Ticks() == o.Ticks(); }
  bool operator>(const DateTime& o) const { return Ticks() > o.Ticks(); }
};

struct Civil { int y, m, d, hh, mm, ss; int64_t frac; };

Civil ToCivil(int64_t ticks) {
  int64_t days = ticks / kTicksPerDay;  // days since 0001-01-01
  int64_t tod = ticks % kTicksPerDay;
  // days since 1970-01-01, then the usual era arithmetic (proleptic Gregorian)
  int64_t z = days - 719162 + 719468;
  int64_t era = (z >= 0 ? z : z - 146096) / 146097;
  int64_t doe = z - era * 146097;
  int64_t yoe = (doe - doe / 1460 + doe / 36524 - doe / 146096) / 365;
  int64_t y = yoe + era * 400;
  int64_t doy = doe - (365 * yoe + yoe / 4 - yoe / 100);
  int64_t mp = (5 * doy + 2) / 153;
  int64_t d = doy - (153 * mp + 2) / 5 + 1;
  int64_t m = mp < 10 ? mp + 3 : mp - 9;
  if (m <= 2) y += 1;
  Civil c;
  c.y = (int)y; c.m = (int)m; c.d = (int)d;
  int64_t secs = tod / kTicksPerSecond;
  c.frac = tod % kTicksPerSecond;
  c.hh = (int)(secs / 3600); c.mm = (int)(secs / 60 % 60); c.ss = (int)(secs % 60);
  return c;
}

// value.ToString("yyyy-MM-ddTHH:mm:ss") — Api/Utilities/DateTimeConverter.cs:26-29.
// Drops the fraction, the Kind and any offset.
std::string FormatConverter(const DateTime& v) {
  Civil c = ToCivil(v.Ticks());
  char buf[40];
  snprintf(buf, sizeof buf, "%04d-%02d-%02dT%02d:%02d:%02d", c.y, c.m, c.d, c.hh, c.mm, c.ss);
  return buf;
}

// System.Text.Json's default DateTime writer (what Dapr's SaveStateAsync<TaskModel> stores,
// Api/Services/TasksStoreManager.cs:35): ISO 8601 extended; ".fffffff" with trailing zeros
// trimmed only when the sub-second part is non-zero; "Z" for Utc, "+hh:mm"/"-hh:mm" for
// Local, nothing for Unspecified.  Samples: docs/aca/05-aca-dapr-pubsubapi/index.md:65-66.
std::string FormatSystemTextJson(const DateTime& v, int local_offset_min) {
  Civil c = ToCivil(v.Ticks());
  char buf[64];
  int n = snprintf(buf, sizeof buf, "%04d-%02d-%02dT%02d:%02d:%02d", c.y, c.m, c.d, c.hh, c.mm, c.ss);
  if (c.frac != 0) {
    char f[16];
    snprintf(f, sizeof f, "%07lld", (long long)c.frac);
    int len = 7;
    while (len > 0 && f[len - 1] == '0') --len;
    buf[n++] = '.';
    memcpy(buf + n, f, len);
    n += len;
  }
  if (v.Kind() == TT_KIND_UTC) {
    buf[n++] = 'Z';
  } else if (v.Kind() >= TT_KIND_LOCAL) {  // 3 = Local with the ambiguous-DST bit
    int off = local_offset_min;
    char sign = off < 0 ? '-' : '+';
    if (off < 0) off = -off;
    n += snprintf(buf + n, sizeof buf - n, "%c%02d:%02d", sign, off / 60, off % 60);
  }
  buf[n] = 0;
  return buf;
}

// ---- System.Guid (Appendix A.5): canonical big-endian bytes <-> "D" string ---------------
struct Guid {
  uint8_t b[16];
  bool Equals(const Guid& o) const { return memcmp(b, o.b, 16) == 0; }
  std::string ToString() const {
    static const char* hex = "0123456789abcdef";
    std::string s;
    for (int i = 0; i < 16; ++i) {
      if (i == 4 || i == 6 || i == 8 || i == 10) s.push_back('-');
      s.push_back(hex[b[i] >> 4]);
      s.push_back(hex[b[i] & 15]);
    }
    return s;
  }
};

std::u16string Utf8ToUtf16(const char* p, int len) {
  std::u16string out;
  int i = 0;
  while (i < len) {
    uint32_t c = (uint8_t)p[i];
    uint32_t cp;
    int extra;
    if (c < 0x80) { cp = c; extra = 0; }
    else if ((c >> 5) == 6) { cp = c & 0x1F; extra = 1; }
    else if ((c >> 4) == 14) { cp = c & 0x0F; extra = 2; }
    else if ((c >> 3) == 30) { cp = c & 0x07; extra = 3; }
    else { cp = 0xFFFD; extra = 0; }
    int j = 0;
    for (; j < extra && i + 1 + j < len && (((uint8_t)p[i + 1 + j]) >> 6) == 2; ++j)
      cp = (cp << 6) | ((uint8_t)p[i + 1 + j] & 0x3F);
    if (j < extra) cp = 0xFFFD;
    i += 1 + j;
    if (cp >= 0x10000) {
      cp -= 0x10000;
      out.push_back((char16_t)(0xD800 + (cp >> 10)));
      out.push_back((char16_t)(0xDC00 + (cp & 0x3FF)));
    } else {
      out.push_back((char16_t)cp);
    }
  }
  return out;
}

// ---- Api/Models/TaskModel.cs:3-13 -------------------------------------------------------
struct TaskModel {
  Guid TaskId;
  std::u16string TaskName;
  std::u16string TaskCreatedBy;
  DateTime TaskCreatedOn;
  DateTime TaskDueDate;
  std::u16string TaskAssignedTo;
  bool IsCompleted = false;
  bool IsOverDue = false;
  // A C# string may hold null: the initialisers (= string.Empty, TaskModel.cs:6,7,10) only apply when a
  // document does not mention the member; `"taskName":null` deserialises to null and serialises back as null.
  // Bit 0 TaskName, bit 1 TaskCreatedBy, bit 2 TaskAssignedTo (the std::u16string is then empty and unused).
  uint8_t null_strings = 0;
  int64_t row = -1;  // insertion index; test bookkeeping, not a reference field
  bool deleted = false;  // bookkeeping: removed from the list / store
};

struct NotImplementedException {};
struct NullReferenceException {};

// Clock inputs the reference reads from the environment, made explicit.
struct Clock {
  int64_t today_local_ticks;  // DateTime.Today (local midnight, Kind Local)
  int64_t run_at_utc_ticks;   // DateTime.UtcNow in the cron handler
  int local_offset_min;       // local UTC offset, only used to print Kind=Local values
  bool string_eq;             // STORE: compare JSON strings (faithful) or ticks/kind (fast)
};

using TaskList = std::vector<TaskModel*>;

// string.Equals(string): ordinal, case-sensitive UTF-16 code-unit equality (Appendix A.4)
inline bool OrdinalEquals(const std::u16string& a, const std::u16string& b) {
  return a.size() == b.size() && memcmp(a.data(), b.data(), a.size() * sizeof(char16_t)) == 0;
}

// ---- Api/Services/FakeTasksManager.cs ---------------------------------------------------
struct FakeTasksManager {
  const TaskList& _tasksList;  // List<TaskModel> _tasksList (:7)

  // :62-66  _tasksList.FirstOrDefault(t => t.TaskId.Equals(taskId))
  TaskModel* GetTaskById(const Guid& taskId) const {
    for (TaskModel* t : _tasksList)
      if (t->TaskId.Equals(taskId)) return t;
    return nullptr;
  }
  // :74-85
  bool MarkTaskCompleted(const Guid& taskId) const {
    TaskModel* task = GetTaskById(taskId);
    if (task != nullptr) { task->IsCompleted = true; return true; }
    return false;
  }
  // :87-100
  bool UpdateTask(const Guid& taskId, const std::u16string& taskName, const std::u16string& assignedTo, DateTime dueDate) const {
    TaskModel* task = GetTaskById(taskId);
    if (task != nullptr) {
      task->TaskName = taskName;
      task->TaskAssignedTo = assignedTo;
      task->TaskDueDate = dueDate;
      return true;
    }
    return false;
  }

  // :107-112  _tasksList.Where(t => t.TaskDueDate.Equals(DateTime.Today.AddDays(-1))).ToList()
  TaskList GetYesterdaysDueTasks(const Clock& clk) const {
    DateTime yesterday = DateTime(clk.today_local_ticks, TT_KIND_LOCAL).AddDays(-1);
    TaskList tasksList;
    for (TaskModel* t : _tasksList)
      if (t->TaskDueDate.Equals(yesterday)) tasksList.push_back(t);
    return tasksList;
  }

  // :102-105
  void MarkOverdueTasks(const TaskList&) const { throw NotImplementedException(); }

  // :68-72  Where(TaskCreatedBy.Equals(createdBy)).OrderByDescending(TaskCreatedOn).ToList()
  TaskList GetTasksByCreator(const std::u16string& createdBy) const {
    TaskList tasksList;
    for (TaskModel* t : _tasksList) {
      if (t->null_strings & 2) throw NullReferenceException();  // t.TaskCreatedBy.Equals(...) on a null string
      if (OrdinalEquals(t->TaskCreatedBy, createdBy)) tasksList.push_back(t);
    }
    // OrderByDescending is a stable sort: equal keys keep source order
    std::stable_sort(tasksList.begin(), tasksList.end(), [](const TaskModel* a, const TaskModel* b) {
      return a->TaskCreatedOn.Ticks() > b->TaskCreatedOn.Ticks();
    });
    return tasksList;
  }
};

// ---- Api/Services/TasksStoreManager.cs --------------------------------------------------
// The Dapr state store is modelled as the documents in insertion order plus a key index
// (key = TaskId.ToString(), :35).  The store-side EQ filter is exact string equality on the
// stored JSON value.
struct TasksStoreManager {
  TaskList* docs;                                   // the stored documents, in store order
  std::unordered_map<std::string, int64_t>* keys;   // key -> insertion index
  TaskList* by_row;                                 // insertion index -> document

  // :47-52  GetStateAsync<TaskModel>(key): the stored document or null
  TaskModel* GetTaskById(const Guid& taskId) const {
    for (TaskModel* t : *docs)
      if (t->TaskId.Equals(taskId)) return t;
    return nullptr;
  }
  // :71-82
  bool MarkTaskCompleted(const Guid& taskId) const {
    TaskModel* taskModel = GetTaskById(taskId);
    if (taskModel != nullptr) { taskModel->IsCompleted = true; return true; }
    return false;
  }
  // :84-102 — `taskModel.TaskAssignedTo` is read before the null check (:88): a missing id throws
  bool UpdateTask(const Guid& taskId, const std::u16string& taskName, const std::u16string& assignedTo, DateTime dueDate) const {
    TaskModel* taskModel = GetTaskById(taskId);
    if (taskModel == nullptr) throw NullReferenceException();
    taskModel->TaskName = taskName;
    taskModel->TaskAssignedTo = assignedTo;
    taskModel->TaskDueDate = dueDate;
    return true;
  }

  static bool EqDueDate(const TaskModel* t, const DateTime& yesterday, const std::string& literal, const Clock& clk) {
    if (clk.string_eq) return FormatSystemTextJson(t->TaskDueDate, clk.local_offset_min) == literal;
    // equivalent arithmetic form: whole-second midnight of that day, Kind Unspecified
    return t->TaskDueDate.Ticks() == yesterday.Ticks() && t->TaskDueDate.Kind() == TT_KIND_UNSPECIFIED;
  }

  // :104-139
  std::vector<TaskModel> GetYesterdaysDueTasks(const Clock& clk) const {
    DateTime yesterday = DateTime(clk.today_local_ticks, TT_KIND_LOCAL).AddDays(-1);  // :119
    std::string jsonDate = FormatConverter(yesterday);                                // :121 (quotes implied)
    std::vector<const TaskModel*> queryResults;                                       // :130 QueryStateAsync
    for (const TaskModel* t : *docs)
      if (EqDueDate(t, yesterday, jsonDate, clk)) queryResults.push_back(t);
    std::vector<TaskModel> tasksList;
    for (const TaskModel* q : queryResults)                                          // :133-135
      if (q->IsCompleted == false && q->IsOverDue == false) tasksList.push_back(*q);  // deserialised copies
    std::stable_sort(tasksList.begin(), tasksList.end(), [](const TaskModel& a, const TaskModel& b) {
      return a.TaskCreatedOn.Ticks() < b.TaskCreatedOn.Ticks();                       // :136 OrderBy, stable
    });
    return tasksList;
  }

  // :141-149  foreach: IsOverDue = true; SaveStateAsync(key = TaskId.ToString(), whole record)
  // Returns the number of keys that were not present (the reference would create them).
  int64_t MarkOverdueTasks(std::vector<TaskModel>& overDueTasksList) const {
    int64_t created = 0;
    for (TaskModel& taskModel : overDueTasksList) {
      taskModel.IsOverDue = true;
      auto it = keys->find(taskModel.TaskId.ToString());
      if (it == keys->end()) { ++created; continue; }  // upsert of an unknown key: counted, not stored
      TaskModel* doc = (*by_row)[it->second];
      int64_t row = doc->row;
      *doc = taskModel;  // the whole record is overwritten with the caller's copy
      doc->row = row;
    }
    return created;
  }

  // :54-69
  std::vector<TaskModel> GetTasksByCreator(const std::u16string& createdBy) const {
    std::vector<TaskModel> tasksList;
    for (const TaskModel* t : *docs) {
      if (t->null_strings & 2) continue;  // the store's EQ on a JSON null matches no string
      if (OrdinalEquals(t->TaskCreatedBy, createdBy)) tasksList.push_back(*t);
    }
    std::stable_sort(tasksList.begin(), tasksList.end(), [](const TaskModel& a, const TaskModel& b) {
      return a.TaskCreatedOn.Ticks() > b.TaskCreatedOn.Ticks();                       // :66 OrderByDescending
    });
    return tasksList;
  }
};

// ---- System.Text.Json: List<TaskModel> as the reference sends it ---------------------------------
// What the controllers return (Api/Controllers/OverdueTasksController.cs:22-26,
// Api/Controllers/TasksController.cs:21-25) is serialised by ASP.NET Core's
// SystemTextJsonOutputFormatter with JsonSerializerDefaults.Web: camelCase names, declaration
// order (Api/Models/TaskModel.cs:5-12), no whitespace, and — because the app configures no
// encoder — JavaScriptEncoder.UnsafeRelaxedJsonEscaping (flavor 0).  Dapr's client
// (SaveStateAsync / PublishEventAsync, Api/Services/TasksStoreManager.cs:35,155) uses the same
// options with the default encoder, JavaScriptEncoder.Default (flavor 1).
// Third-party code absent from /root/reference: System.Text.Json / System.Text.Encodings.Web of
// .NET 9 (global.json:3), Dapr.Client 1.14.0.  Their published behaviour, restated:
//   * both encoders write \b \t \n \f \r as two characters, every other code unit they refuse as
//     \uXXXX with upper-case hex, a supplementary-plane scalar as its two surrogates;
//   * Default allows only Basic Latin (U+0020..U+007E) minus  " & ' + < > ` \  ; '\' is "\\", the
//     quote is ";
//   * UnsafeRelaxed writes \" and \\, and allows every other BMP code point that the encoder
//     treats as defined (TextEncoderSettings.ForbidUndefinedCharacters): general category letter,
//     mark, number, punctuation or symbol, or U+0020 — so controls (Cc), format characters (Cf),
//     the other separators (Zs, Zl, Zp), surrogates, private use and unassigned code points are
//     escaped.  The table is include/tt_bmp_defined.h, generated from Python's unicodedata
//     (scripts/gen_unicode_bitmap.py; its Unicode version is older than the one .NET 9 carries,
//     code points assigned in between differ — the one remaining assumption for such text).
// Pins: the documented sample body docs/aca/05-aca-dapr-pubsubapi/index.md:60-69 (tests/test_oracle.py).
bool JsonAllowed(char16_t c, int flavor) {
  if (c < 0x20 || c == u'"' || c == u'\\' || c == 0x7F) return false;
  if (c < 0x7F) {
    if (flavor == 1) return !(c == u'&' || c == u'\'' || c == u'+' || c == u'<' || c == u'>' || c == u'`');
    return true;
  }
  if (flavor == 1) return false;  // outside UnicodeRanges.BasicLatin
  static const uint32_t kDefined[2048] = {TT_BMP_DEFINED_WORDS};
  return (kDefined[c >> 5] >> (c & 31)) & 1u;
}

void JsonWriteString(std::string& out, const std::u16string& s, int flavor) {
  static const char* HEX = "0123456789ABCDEF";
  out.push_back('"');
  for (char16_t c : s) {
    if (JsonAllowed(c, flavor)) {  // UTF-8 of a BMP code unit
      if (c < 0x80) out.push_back((char)c);
      else if (c < 0x800) { out.push_back((char)(0xC0 | (c >> 6))); out.push_back((char)(0x80 | (c & 0x3F))); }
      else { out.push_back((char)(0xE0 | (c >> 12))); out.push_back((char)(0x80 | ((c >> 6) & 0x3F))); out.push_back((char)(0x80 | (c & 0x3F))); }
      continue;
    }
    switch (c) {
      case u'\b': out += "\\b"; continue;
      case u'\t': out += "\\t"; continue;
      case u'\n': out += "\\n"; continue;
      case u'\f': out += "\\f"; continue;
      case u'\r': out += "\\r"; continue;
      case u'\\': out += "\\\\"; continue;
      case u'"': if (flavor == 0) { out += "\\\""; continue; } break;
      default: break;
    }
    out += "\\u";
    out.push_back(HEX[(c >> 12) & 15]);
    out.push_back(HEX[(c >> 8) & 15]);
    out.push_back(HEX[(c >> 4) & 15]);
    out.push_back(HEX[c & 15]);
  }
  out.push_back('"');
}

void JsonWriteTaskModel(std::string& out, const TaskModel& t, int flavor, int local_offset_min) {
  out += "{\"taskId\":\"";
  out += t.TaskId.ToString();
  auto str = [&](const std::u16string& v, int bit) {
    if (t.null_strings & bit) out += "null";  // JsonSerializer writes a null string as the literal
    else JsonWriteString(out, v, flavor);
  };
  out += "\",\"taskName\":";
  str(t.TaskName, 1);
  out += ",\"taskCreatedBy\":";
  str(t.TaskCreatedBy, 2);
  out += ",\"taskCreatedOn\":\"";
  out += FormatSystemTextJson(t.TaskCreatedOn, local_offset_min);
  out += "\",\"taskDueDate\":\"";
  out += FormatSystemTextJson(t.TaskDueDate, local_offset_min);
  out += "\",\"taskAssignedTo\":";
  str(t.TaskAssignedTo, 4);
  out += t.IsCompleted ? ",\"isCompleted\":true" : ",\"isCompleted\":false";
  out += t.IsOverDue ? ",\"isOverDue\":true" : ",\"isOverDue\":false";
  out.push_back('}');
}

// ---- System.Text.Json: one stored document back into a TaskModel --------------------------------
// What GetStateAsync<TaskModel> / QueryStateAsync<TaskModel> (Api/Services/TasksStoreManager.cs:50,61,130)
// and the queue handler's [FromBody] TaskModel (Svc/Controllers/ExternalTasksProcessorController.cs:29)
// do with a document: JsonSerializer.Deserialize<TaskModel> with JsonSerializerDefaults.Web.
// Restated behaviour of System.Text.Json (.NET 9, not under /root/reference):
//   * the text must be one JSON value; here it has to be an object (anything else throws);
//   * property names are matched ignoring case (PropertyNameCaseInsensitive), unknown ones are
//     skipped, a repeated one overwrites, a missing one leaves the initialiser's value;
//   * Guid: exactly the 36-character "D" form; bool: the literals; string: any JSON string or null;
//   * DateTime: ISO 8601-1 extended profile — YYYY-MM-DD, optionally 'T'hh:mm[:ss[.fraction]] and
//     'Z' (Kind Utc) or (+|-)hh[:mm] (the instant in local time, Kind Local), else Kind Unspecified;
//     1..16 fraction digits, those beyond 100 ns dropped;
//   * lone surrogate escapes, raw control characters in strings, trailing commas and comments throw.
// This restatement builds a small value tree first and maps its members afterwards (the product
// streams through the text once); a null string member stays null (TaskModel::null_strings).
struct JsonNode {
  enum Type { Null, Bool, Number, String, Array, Object } type = Null;
  bool b = false;
  std::string text;  // String: unescaped UTF-8; Number: the literal
  std::vector<std::pair<std::string, JsonNode>> members;
  std::vector<JsonNode> items;
};

struct JsonParseError {};

struct JsonText {
  const char* p;
  const char* end;
  void Ws() { while (p < end && (*p == ' ' || *p == '\t' || *p == '\n' || *p == '\r')) ++p; }
  char Peek() { if (p >= end) throw JsonParseError(); return *p; }
  void Expect(char c) { if (p >= end || *p != c) throw JsonParseError(); ++p; }
  static int Hex(char c) {
    if (c >= '0' && c <= '9') return c - '0';
    if (c >= 'a' && c <= 'f') return c - 'a' + 10;
    if (c >= 'A' && c <= 'F') return c - 'A' + 10;
    throw JsonParseError();
  }
  uint32_t Unit() {
    if (end - p < 4) throw JsonParseError();
    uint32_t v = 0;
    for (int i = 0; i < 4; ++i) v = v * 16 + Hex(p[i]);
    p += 4;
    return v;
  }
  static void Append(std::string& s, uint32_t cp) {
    if (cp < 0x80) s.push_back((char)cp);
    else if (cp < 0x800) { s.push_back((char)(0xC0 | (cp >> 6))); s.push_back((char)(0x80 | (cp & 63))); }
    else if (cp < 0x10000) { s.push_back((char)(0xE0 | (cp >> 12))); s.push_back((char)(0x80 | ((cp >> 6) & 63))); s.push_back((char)(0x80 | (cp & 63))); }
    else { s.push_back((char)(0xF0 | (cp >> 18))); s.push_back((char)(0x80 | ((cp >> 12) & 63))); s.push_back((char)(0x80 | ((cp >> 6) & 63))); s.push_back((char)(0x80 | (cp & 63))); }
  }
  std::string Str() {
    Expect('"');
    std::string s;
    while (true) {
      if (p >= end) throw JsonParseError();
      unsigned char c = (unsigned char)*p++;
      if (c == '"') return s;
      if (c < 0x20) throw JsonParseError();
      if (c == '\\') {
        if (p >= end) throw JsonParseError();
        char e = *p++;
        switch (e) {
          case '"': s.push_back('"'); break;
          case '\\': s.push_back('\\'); break;
          case '/': s.push_back('/'); break;
          case 'b': s.push_back('\b'); break;
          case 'f': s.push_back('\f'); break;
          case 'n': s.push_back('\n'); break;
          case 'r': s.push_back('\r'); break;
          case 't': s.push_back('\t'); break;
          case 'u': {
            uint32_t u = Unit();
            if (u >= 0xDC00 && u <= 0xDFFF) throw JsonParseError();
            if (u >= 0xD800 && u <= 0xDBFF) {
              if (end - p < 2 || p[0] != '\\' || p[1] != 'u') throw JsonParseError();
              p += 2;
              uint32_t lo = Unit();
              if (lo < 0xDC00 || lo > 0xDFFF) throw JsonParseError();
              u = 0x10000 + ((u - 0xD800) << 10) + (lo - 0xDC00);
            }
            Append(s, u);
            break;
          }
          default: throw JsonParseError();
        }
        continue;
      }
      if (c >= 0x80) {  // a well-formed multi-byte sequence
        int extra = (c >> 5) == 6 ? 1 : (c >> 4) == 14 ? 2 : (c >> 3) == 30 ? 3 : -1;
        if (extra < 0 || end - p < extra) throw JsonParseError();
        s.push_back((char)c);
        for (int i = 0; i < extra; ++i) {
          unsigned char d = (unsigned char)*p++;
          if ((d >> 6) != 2) throw JsonParseError();
          s.push_back((char)d);
        }
        continue;
      }
      s.push_back((char)c);
    }
  }
  JsonNode Value() {
    Ws();
    JsonNode n;
    char c = Peek();
    if (c == '{') {
      ++p;
      n.type = JsonNode::Object;
      Ws();
      if (Peek() == '}') { ++p; return n; }
      while (true) {
        Ws();
        std::string k = Str();
        Ws();
        Expect(':');
        n.members.emplace_back(std::move(k), Value());
        Ws();
        if (Peek() == ',') { ++p; continue; }
        Expect('}');
        return n;
      }
    }
    if (c == '[') {
      ++p;
      n.type = JsonNode::Array;
      Ws();
      if (Peek() == ']') { ++p; return n; }
      while (true) {
        n.items.push_back(Value());
        Ws();
        if (Peek() == ',') { ++p; continue; }
        Expect(']');
        return n;
      }
    }
    if (c == '"') { n.type = JsonNode::String; n.text = Str(); return n; }
    auto word = [&](const char* w) { size_t l = strlen(w); if ((size_t)(end - p) < l || memcmp(p, w, l) != 0) throw JsonParseError(); p += l; };
    if (c == 't') { word("true"); n.type = JsonNode::Bool; n.b = true; return n; }
    if (c == 'f') { word("false"); n.type = JsonNode::Bool; n.b = false; return n; }
    if (c == 'n') { word("null"); n.type = JsonNode::Null; return n; }
    if (c == '-' || (c >= '0' && c <= '9')) {
      n.type = JsonNode::Number;
      while (p < end && (*p == '-' || *p == '+' || *p == '.' || *p == 'e' || *p == 'E' || (*p >= '0' && *p <= '9'))) n.text.push_back(*p++);
      return n;
    }
    throw JsonParseError();
  }
};

bool IEquals(const std::string& a, const char* b) {
  size_t n = strlen(b);
  if (a.size() != n) return false;
  for (size_t i = 0; i < n; ++i) {
    char x = a[i], y = b[i];
    if (x >= 'A' && x <= 'Z') x = (char)(x - 'A' + 'a');
    if (y >= 'A' && y <= 'Z') y = (char)(y - 'A' + 'a');
    if (x != y) return false;
  }
  return true;
}

int Digits(const std::string& s, size_t at, int n) {
  if (at + n > s.size()) throw JsonParseError();
  int v = 0;
  for (int i = 0; i < n; ++i) {
    char c = s[at + i];
    if (c < '0' || c > '9') throw JsonParseError();
    v = v * 10 + (c - '0');
  }
  return v;
}

// Utf8JsonReader.GetDateTime
DateTime JsonReadDateTime(const std::string& s, int local_offset_min) {
  static const int kCum[12] = {0, 31, 59, 90, 120, 151, 181, 212, 243, 273, 304, 334};
  if (s.size() < 10 || s.size() > 48) throw JsonParseError();
  int y = Digits(s, 0, 4), mo = Digits(s, 5, 2), d = Digits(s, 8, 2);
  if (s[4] != '-' || s[7] != '-' || y < 1 || mo < 1 || mo > 12 || d < 1) throw JsonParseError();
  bool leap = (y % 4 == 0 && y % 100 != 0) || y % 400 == 0;
  int dim = (mo == 2) ? (leap ? 29 : 28) : ((mo == 4 || mo == 6 || mo == 9 || mo == 11) ? 30 : 31);
  if (d > dim) throw JsonParseError();
  int64_t py = y - 1;
  int64_t days = py * 365 + py / 4 - py / 100 + py / 400 + kCum[mo - 1] + ((leap && mo > 2) ? 1 : 0) + (d - 1);
  int64_t ticks = days * kTicksPerDay;
  uint32_t kind = TT_KIND_UNSPECIFIED;
  size_t i = 10;
  if (i < s.size()) {
    if (s[i] != 'T') throw JsonParseError();
    int hh = Digits(s, 11, 2), mi = Digits(s, 14, 2), ss = 0;
    if (s[13] != ':' || hh > 23 || mi > 59) throw JsonParseError();
    int64_t frac = 0;
    i = 16;
    if (i < s.size() && s[i] == ':') {
      ss = Digits(s, 17, 2);
      if (ss > 59) throw JsonParseError();
      i = 19;
      if (i < s.size() && (s[i] == '.' || s[i] == ',')) {
        ++i;
        int nd = 0;
        while (i < s.size() && s[i] >= '0' && s[i] <= '9') {
          if (nd < 7) frac = frac * 10 + (s[i] - '0');
          ++nd;
          ++i;
        }
        if (nd == 0 || nd > 16) throw JsonParseError();
        while (nd++ < 7) frac *= 10;
      }
    }
    ticks += ((int64_t)hh * 3600 + mi * 60 + ss) * kTicksPerSecond + frac;
    if (i < s.size()) {
      if (s[i] == 'Z') {
        kind = TT_KIND_UTC;
        ++i;
      } else if (s[i] == '+' || s[i] == '-') {
        int sign = s[i] == '-' ? -1 : 1;
        int oh = Digits(s, i + 1, 2), om = 0;
        i += 3;
        if (i < s.size()) {
          if (s[i] != ':') throw JsonParseError();
          om = Digits(s, i + 1, 2);
          i += 3;
        }
        if (oh > 14 || om > 59) throw JsonParseError();
        // DateTimeOffset(...).LocalDateTime: the same instant on the local clock
        ticks += (int64_t)(local_offset_min - sign * (oh * 60 + om)) * 60 * kTicksPerSecond;
        kind = TT_KIND_LOCAL;
      }
      if (i != s.size()) throw JsonParseError();
    }
  }
  if (ticks < 0 || ticks > 3155378975999999999LL) throw JsonParseError();
  return DateTime(ticks, kind);
}

Guid JsonReadGuid(const std::string& s) {
  if (s.size() != 36) throw JsonParseError();
  Guid g;
  int b = 0;
  for (size_t i = 0; i < 36;) {
    if (i == 8 || i == 13 || i == 18 || i == 23) {
      if (s[i] != '-') throw JsonParseError();
      ++i;
      continue;
    }
    g.b[b++] = (uint8_t)(JsonText::Hex(s[i]) * 16 + JsonText::Hex(s[i + 1]));
    i += 2;
  }
  return g;
}

// JsonSerializer.Deserialize<TaskModel>(text, web options); throws JsonParseError like JsonException
TaskModel DeserializeTaskModel(const char* p, const char* end, int local_offset_min) {
  JsonText t{p, end};
  JsonNode root = t.Value();
  t.Ws();
  if (t.p != end || root.type != JsonNode::Object) throw JsonParseError();
  TaskModel m;  // initialisers: Guid.Empty, string.Empty, default(DateTime), false
  memset(m.TaskId.b, 0, 16);
  auto str = [&m](const JsonNode& v, uint8_t bit) {
    m.null_strings = (uint8_t)(m.null_strings & ~bit);  // a repeated member overwrites, null-ness included
    if (v.type == JsonNode::Null) {
      m.null_strings |= bit;
      return std::u16string();
    }
    if (v.type != JsonNode::String) throw JsonParseError();
    return Utf8ToUtf16(v.text.data(), (int)v.text.size());
  };
  for (auto& kv : root.members) {
    const std::string& k = kv.first;
    const JsonNode& v = kv.second;
    if (IEquals(k, "taskId")) { if (v.type != JsonNode::String) throw JsonParseError(); m.TaskId = JsonReadGuid(v.text); }
    else if (IEquals(k, "taskName")) m.TaskName = str(v, 1);
    else if (IEquals(k, "taskCreatedBy")) m.TaskCreatedBy = str(v, 2);
    else if (IEquals(k, "taskAssignedTo")) m.TaskAssignedTo = str(v, 4);
    else if (IEquals(k, "taskCreatedOn")) { if (v.type != JsonNode::String) throw JsonParseError(); m.TaskCreatedOn = JsonReadDateTime(v.text, local_offset_min); }
    else if (IEquals(k, "taskDueDate")) { if (v.type != JsonNode::String) throw JsonParseError(); m.TaskDueDate = JsonReadDateTime(v.text, local_offset_min); }
    else if (IEquals(k, "isCompleted")) { if (v.type != JsonNode::Bool) throw JsonParseError(); m.IsCompleted = v.b; }
    else if (IEquals(k, "isOverDue")) { if (v.type != JsonNode::Bool) throw JsonParseError(); m.IsOverDue = v.b; }
  }
  return m;
}

struct OracleTable {
  TaskList rows;  // every task ever added, by insertion index; owns the TaskModel objects
  TaskList live;  // the reference's _tasksList / the store's documents, in list order
  std::unordered_map<std::string, int64_t> keys;
  bool keys_built = false;
  ~OracleTable() { for (TaskModel* t : rows) delete t; }
  void BuildKeys() {
    if (keys_built) return;
    keys.clear();
    keys.reserve(live.size() * 2);
    for (TaskModel* t : live) keys[t->TaskId.ToString()] = t->row;
    keys_built = true;
  }
  void Add(TaskModel* t) { t->row = (int64_t)rows.size(); rows.push_back(t); live.push_back(t); keys_built = false; }
  // List<T>.Remove(item) / DeleteStateAsync(key): later items keep their relative order
  void Remove(TaskModel* t) {
    live.erase(std::find(live.begin(), live.end(), t));
    t->deleted = true;
    keys_built = false;
  }
};

void GuidFromHiLo(uint64_t hi, uint64_t lo, Guid* g) {
  for (int i = 0; i < 8; ++i) g->b[i] = (uint8_t)(hi >> (56 - 8 * i));
  for (int i = 0; i < 8; ++i) g->b[8 + i] = (uint8_t)(lo >> (56 - 8 * i));
}

TaskModel* SynthTask(const tt_synth_cfg* cfg, uint64_t grow) {
  tt_synth_row r = tt_synth_make(cfg, grow);
  TaskModel* t = new TaskModel();
  GuidFromHiLo(r.id_hi, r.id_lo, &t->TaskId);
  char buf[48];
  int n = tt_synth_creator(r.creator, buf);
  t->TaskCreatedBy = Utf8ToUtf16(buf, n);
  char name[64];
  int nn = tt_synth_name(tt_synth_name_ref(cfg, grow), name);
  t->TaskName = Utf8ToUtf16(name, nn);
  nn = tt_synth_assignee(tt_synth_assignee_ref(cfg, grow), name);
  t->TaskAssignedTo = Utf8ToUtf16(name, nn);
  t->TaskCreatedOn = DateTime(r.created_raw);
  t->TaskDueDate = DateTime(r.due_raw);
  t->IsCompleted = r.completed;
  t->IsOverDue = r.overdue;
  return t;
}

}  // namespace

extern "C" {

typedef struct OracleTable orc_table;

#define ORC_MODE_FAKE 0
#define ORC_MODE_STORE 1

orc_table* orc_create(void) { return new OracleTable(); }
void orc_destroy(orc_table* t) { delete t; }
int64_t orc_count(orc_table* t) { return (int64_t)t->rows.size(); }

int64_t orc_add(orc_table* t, const uint8_t* id16, const char* name, int32_t name_len, const char* created_by,
                int32_t created_by_len, const char* assigned, int32_t assigned_len, uint64_t created_raw,
                uint64_t due_raw, int32_t completed, int32_t overdue) {
  TaskModel* m = new TaskModel();
  memcpy(m->TaskId.b, id16, 16);
  m->TaskName = Utf8ToUtf16(name, name_len);
  m->TaskCreatedBy = Utf8ToUtf16(created_by, created_by_len);
  m->TaskAssignedTo = Utf8ToUtf16(assigned, assigned_len);
  m->TaskCreatedOn = DateTime(created_raw);
  m->TaskDueDate = DateTime(due_raw);
  m->IsCompleted = completed != 0;
  m->IsOverDue = overdue != 0;
  t->Add(m);
  return m->row;
}

int64_t orc_add_synth(orc_table* t, const tt_synth_cfg* cfg, uint64_t first_row, uint64_t n) {
  t->rows.reserve(t->rows.size() + n);
  for (uint64_t i = 0; i < n; ++i) t->Add(SynthTask(cfg, first_row + i));
  return (int64_t)t->rows.size();
}

// Api/Services/FakeTasksManager.cs:10-25 with the clock and the RNG injected.  TaskIds are
// v4-shaped values from a splitmix stream (Guid.NewGuid() is random in the reference).
void orc_seed_fake(orc_table* t, int64_t now_utc_ticks, uint64_t rng_seed) {
  DateTime now(now_utc_ticks, TT_KIND_UTC);
  for (int i = 0; i < 10; ++i) {
    TaskModel* m = new TaskModel();
    uint64_t hi = tt_mix64(rng_seed + 2 * i), lo = tt_mix64(rng_seed + 2 * i + 1);
    hi = (hi & ~0xF000ULL) | 0x4000ULL;
    lo = (lo & 0x3FFFFFFFFFFFFFFFULL) | 0x8000000000000000ULL;
    GuidFromHiLo(hi, lo, &m->TaskId);
    char buf[64];
    int n = snprintf(buf, sizeof buf, "Task number: %d", i);
    m->TaskName = Utf8ToUtf16(buf, n);
    const char* by = "tjoudeh@bitoftech.net";
    m->TaskCreatedBy = Utf8ToUtf16(by, (int)strlen(by));
    m->TaskCreatedOn = now.AddMinutes(i);
    m->TaskDueDate = now.AddDays(i);
    n = snprintf(buf, sizeof buf, "assignee%d@mail.com", (int)(tt_mix64(rng_seed + 100 + i) % 50));
    m->TaskAssignedTo = Utf8ToUtf16(buf, n);
    t->Add(m);
  }
}

static Clock MakeClock(int64_t today, int64_t run_at, int32_t off, int32_t string_eq) {
  Clock c;
  c.today_local_ticks = today;
  c.run_at_utc_ticks = run_at;
  c.local_offset_min = off;
  c.string_eq = string_eq != 0;
  return c;
}

static int64_t Emit(const std::vector<int64_t>& rows, uint32_t* out, int64_t cap) {
  for (int64_t i = 0; i < (int64_t)rows.size() && i < cap; ++i) out[i] = (uint32_t)rows[i];
  return (int64_t)rows.size();
}

// ITasksManager.GetYesterdaysDueTasks — returns the count; writes up to cap row numbers in
// the order the reference's list has.
int64_t orc_get_yesterdays_due(orc_table* t, int32_t mode, int64_t today_local_ticks, int32_t local_offset_min,
                               int32_t string_eq, uint32_t* out_rows, int64_t cap) {
  Clock clk = MakeClock(today_local_ticks, 0, local_offset_min, string_eq);
  std::vector<int64_t> rows;
  if (mode == ORC_MODE_FAKE) {
    FakeTasksManager m{t->live};
    for (TaskModel* x : m.GetYesterdaysDueTasks(clk)) rows.push_back(x->row);
  } else {
    TasksStoreManager m{&t->live, &t->keys, &t->rows};
    for (const TaskModel& x : m.GetYesterdaysDueTasks(clk)) rows.push_back(x.row);
  }
  return Emit(rows, out_rows, cap);
}

// Svc/Controllers/ScheduledTasksManagerController.cs:19-46 — the whole cron tick:
// GET api/overduetasks -> keep runAt.Date > TaskDueDate.Date -> POST markoverdue if any.
// Returns 0, or 1 when the manager threw NotImplementedException (FAKE mode with survivors).
int32_t orc_cron_job(orc_table* t, int32_t mode, int64_t today_local_ticks, int64_t run_at_utc_ticks,
                     int32_t local_offset_min, int32_t string_eq, int64_t* n_listed, int64_t* n_overdue,
                     uint32_t* out_rows, int64_t cap) {
  Clock clk = MakeClock(today_local_ticks, run_at_utc_ticks, local_offset_min, string_eq);
  DateTime runAt(clk.run_at_utc_ticks, TT_KIND_UTC);  // :22
  std::vector<int64_t> listed;
  int32_t status = 0;
  if (mode == ORC_MODE_FAKE) {
    FakeTasksManager m{t->live};
    TaskList tasksList = m.GetYesterdaysDueTasks(clk);  // :28
    TaskList overdueTasksList;
    for (TaskModel* taskModel : tasksList) {            // :32-38
      listed.push_back(taskModel->row);
      if (runAt.Date() > taskModel->TaskDueDate.Date()) overdueTasksList.push_back(taskModel);
    }
    *n_overdue = (int64_t)overdueTasksList.size();
    if (overdueTasksList.size() > 0) {                  // :40-45
      try { m.MarkOverdueTasks(overdueTasksList); } catch (const NotImplementedException&) { status = 1; }
    }
  } else {
    t->BuildKeys();
    TasksStoreManager m{&t->live, &t->keys, &t->rows};
    std::vector<TaskModel> tasksList = m.GetYesterdaysDueTasks(clk);
    std::vector<TaskModel> overdueTasksList;
    for (TaskModel& taskModel : tasksList) {
      listed.push_back(taskModel.row);
      if (runAt.Date() > taskModel.TaskDueDate.Date()) overdueTasksList.push_back(taskModel);
    }
    *n_overdue = (int64_t)overdueTasksList.size();
    if (overdueTasksList.size() > 0) m.MarkOverdueTasks(overdueTasksList);
  }
  *n_listed = (int64_t)listed.size();
  Emit(listed, out_rows, cap);
  return status;
}

// ---- EXTENSION (no reference counterpart) ------------------------------------------
// The generalised predicate: "due yesterday or earlier" instead of "due exactly yesterday at
// 00:00:00".  Written as the two managers would write it, one LINQ clause changed —
//   FAKE : _tasksList.Where(t => t.TaskDueDate.Date <= DateTime.Today.AddDays(-1).Date).ToList()
//   STORE: the same date test in place of the store-side EQ, then :133-136 unchanged
//          (IsCompleted == false && IsOverDue == false, OrderBy(TaskCreatedOn))
// — followed by the unchanged cron handler (ScheduledTasksManagerController.cs:28-45).  It covers what
// docs/aca/07-aca-cron-bindings/index.md:106 calls the edge cases of the simple implementation.
int64_t orc_backlog(orc_table* t, int32_t mode, int32_t mark, int64_t today_local_ticks, int64_t run_at_utc_ticks,
                    int64_t* n_overdue, int32_t* threw, uint32_t* out_rows, int64_t cap) {
  const DateTime yesterday = DateTime(today_local_ticks, TT_KIND_LOCAL).AddDays(-1).Date();
  const DateTime runAt(run_at_utc_ticks, TT_KIND_UTC);
  std::vector<TaskModel*> list;
  for (TaskModel* x : t->live) {
    if (x->TaskDueDate.Date() > yesterday) continue;
    if (mode == ORC_MODE_STORE && !(x->IsCompleted == false && x->IsOverDue == false)) continue;
    list.push_back(x);
  }
  if (mode == ORC_MODE_STORE)
    std::stable_sort(list.begin(), list.end(),
                     [](const TaskModel* a, const TaskModel* b) { return a->TaskCreatedOn.Ticks() < b->TaskCreatedOn.Ticks(); });
  std::vector<int64_t> rows;
  std::vector<TaskModel*> overdue;
  for (TaskModel* x : list) {
    rows.push_back(x->row);
    if (runAt.Date() > x->TaskDueDate.Date()) overdue.push_back(x);
  }
  *n_overdue = (int64_t)overdue.size();
  *threw = 0;
  if (mark && !overdue.empty()) {
    if (mode == ORC_MODE_FAKE) *threw = 1;  // FakeTasksManager.MarkOverdueTasks throws (:104)
    else
      for (TaskModel* x : overdue) x->IsOverDue = true;  // TasksStoreManager.cs:146
  }
  return Emit(rows, out_rows, cap);
}

// ITasksManager.MarkOverdueTasks(list) addressed by TaskId.  Returns 0, or 1 if it threw
// (FAKE).  *n_missing = ids with no stored record.
int32_t orc_mark_overdue(orc_table* t, int32_t mode, const uint8_t* ids16, int64_t n, int64_t* n_missing) {
  *n_missing = 0;
  if (mode == ORC_MODE_FAKE) {
    try { FakeTasksManager{t->live}.MarkOverdueTasks(TaskList()); } catch (const NotImplementedException&) { return 1; }
    return 0;
  }
  t->BuildKeys();
  TasksStoreManager m{&t->live, &t->keys, &t->rows};
  std::vector<TaskModel> list;
  for (int64_t i = 0; i < n; ++i) {
    Guid g;
    memcpy(g.b, ids16 + 16 * i, 16);
    auto it = t->keys.find(g.ToString());
    if (it != t->keys.end()) {
      list.push_back(*t->rows[it->second]);  // the caller's copy of the stored record
    } else {
      TaskModel x;
      x.TaskId = g;
      list.push_back(x);
    }
  }
  *n_missing = m.MarkOverdueTasks(list);
  return 0;
}

int64_t orc_get_by_creator(orc_table* t, int32_t mode, const char* utf8, int32_t len, uint32_t* out_rows, int64_t cap) {
  std::u16string createdBy = Utf8ToUtf16(utf8, len);
  std::vector<int64_t> rows;
  if (mode == ORC_MODE_FAKE) {
    FakeTasksManager m{t->live};
    try {
      for (TaskModel* x : m.GetTasksByCreator(createdBy)) rows.push_back(x->row);
    } catch (const NullReferenceException&) {
      return -2;  // a task whose TaskCreatedBy is null: the in-memory manager's lambda throws
    }
  } else {
    TasksStoreManager m{&t->live, &t->keys, &t->rows};
    for (const TaskModel& x : m.GetTasksByCreator(createdBy)) rows.push_back(x.row);
  }
  return Emit(rows, out_rows, cap);
}

void orc_get_flags(orc_table* t, uint8_t* completed, uint8_t* overdue) {
  for (size_t i = 0; i < t->rows.size(); ++i) {
    completed[i] = !t->rows[i]->deleted && t->rows[i]->IsCompleted;
    overdue[i] = !t->rows[i]->deleted && t->rows[i]->IsOverDue;
  }
}

int64_t orc_live_count(orc_table* t) { return (int64_t)t->live.size(); }

// ---- ITasksManager point operations.  mode selects the implementation; return values:
// get_by_id -> row or -1; mark_completed / update / delete -> 1 true, 0 false, 2 = the reference
// throws (TasksStoreManager.UpdateTask on a missing id: NullReferenceException, :88).
int64_t orc_get_by_id(orc_table* t, int32_t mode, const uint8_t* id16) {
  Guid g;
  memcpy(g.b, id16, 16);
  TaskModel* m = mode == ORC_MODE_FAKE ? FakeTasksManager{t->live}.GetTaskById(g)
                                       : TasksStoreManager{&t->live, &t->keys, &t->rows}.GetTaskById(g);
  return m ? m->row : -1;
}

int32_t orc_mark_completed(orc_table* t, int32_t mode, const uint8_t* id16) {
  Guid g;
  memcpy(g.b, id16, 16);
  return mode == ORC_MODE_FAKE ? FakeTasksManager{t->live}.MarkTaskCompleted(g)
                               : TasksStoreManager{&t->live, &t->keys, &t->rows}.MarkTaskCompleted(g);
}

int32_t orc_update_task(orc_table* t, int32_t mode, const uint8_t* id16, const char* name, int32_t name_len,
                        const char* assigned, int32_t assigned_len, uint64_t due_raw) {
  Guid g;
  memcpy(g.b, id16, 16);
  std::u16string n = Utf8ToUtf16(name, name_len), a = Utf8ToUtf16(assigned, assigned_len);
  if (mode == ORC_MODE_FAKE) return FakeTasksManager{t->live}.UpdateTask(g, n, a, DateTime(due_raw));
  try {
    return TasksStoreManager{&t->live, &t->keys, &t->rows}.UpdateTask(g, n, a, DateTime(due_raw));
  } catch (const NullReferenceException&) {
    return 2;
  }
}

// FakeTasksManager.cs:49-60 (FirstOrDefault + Remove -> true/false);
// TasksStoreManager.cs:40-45 (DeleteStateAsync, returns true whether or not the key existed)
int32_t orc_delete_task(orc_table* t, int32_t mode, const uint8_t* id16) {
  Guid g;
  memcpy(g.b, id16, 16);
  TaskModel* m = FakeTasksManager{t->live}.GetTaskById(g);
  if (m) t->Remove(m);
  return mode == ORC_MODE_FAKE ? (m != nullptr) : 1;
}

// The state store's key index exists before the cron tick runs; build it outside timed regions.
void orc_build_keys(orc_table* t) { t->BuildKeys(); }

// CreateNewTask (FakeTasksManager.cs:32-47 / TasksStoreManager.cs:22-38) is orc_add with the Guid and
// the clock supplied by the caller.

void orc_set_all_flags(orc_table* t, const uint8_t* completed, const uint8_t* overdue) {
  for (size_t i = 0; i < t->rows.size(); ++i) {
    t->rows[i]->IsCompleted = completed[i] != 0;
    t->rows[i]->IsOverDue = overdue[i] != 0;
  }
}

void orc_set_flags(orc_table* t, int64_t row, int32_t completed, int32_t overdue) {
  t->rows[row]->IsCompleted = completed != 0;
  t->rows[row]->IsOverDue = overdue != 0;
}

// Column export so tests can load the same rows into the device table.
void orc_export(orc_table* t, uint8_t* ids16, uint64_t* created_raw, uint64_t* due_raw) {
  for (size_t i = 0; i < t->rows.size(); ++i) {
    memcpy(ids16 + 16 * i, t->rows[i]->TaskId.b, 16);
    created_raw[i] = t->rows[i]->TaskCreatedOn.dateData;
    due_raw[i] = t->rows[i]->TaskDueDate.dateData;
  }
}

// UTF-8 of TaskCreatedBy for one row (ASCII / BMP only is enough for the tests). Returns length.
int32_t orc_created_by(orc_table* t, int64_t row, char* out, int32_t cap) {
  const std::u16string& s = t->rows[row]->TaskCreatedBy;
  int32_t n = 0;
  for (char16_t c : s) {
    if (c < 0x80) { if (n < cap) out[n] = (char)c; n += 1; }
    else if (c < 0x800) { if (n + 1 < cap) { out[n] = (char)(0xC0 | (c >> 6)); out[n + 1] = (char)(0x80 | (c & 0x3F)); } n += 2; }
    else { if (n + 2 < cap) { out[n] = (char)(0xE0 | (c >> 12)); out[n + 1] = (char)(0x80 | ((c >> 6) & 0x3F)); out[n + 2] = (char)(0x80 | (c & 0x3F)); } n += 3; }
  }
  return n;
}

int32_t orc_format_stj(uint64_t raw, int32_t local_offset_min, char* out, int32_t cap) {
  std::string s = FormatSystemTextJson(DateTime(raw), local_offset_min);
  snprintf(out, cap, "%s", s.c_str());
  return (int32_t)s.size();
}

int32_t orc_format_converter(uint64_t raw, char* out, int32_t cap) {
  std::string s = FormatConverter(DateTime(raw));
  snprintf(out, cap, "%s", s.c_str());
  return (int32_t)s.size();
}

// JsonSerializer.Serialize(List<TaskModel>) for the listed rows (insertion indices), in the given
// order.  Returns the length; writes at most cap bytes.
int64_t orc_serialize(orc_table* t, const uint32_t* rows, int64_t n, int32_t flavor, int32_t local_offset_min, char* out,
                      int64_t cap) {
  std::string s = "[";
  for (int64_t i = 0; i < n; ++i) {
    if (i) s.push_back(',');
    JsonWriteTaskModel(s, *t->rows[rows[i]], flavor, local_offset_min);
  }
  s.push_back(']');
  if (out && cap > 0) memcpy(out, s.data(), (size_t)std::min<int64_t>(cap, (int64_t)s.size()));
  return (int64_t)s.size();
}

// UTF-8 of TaskName / TaskAssignedTo for one row (test bookkeeping: lets the tests load the same
// strings into the device table).  which: 0 = TaskName, 2 = TaskAssignedTo.  Returns the length.
int32_t orc_string_field(orc_table* t, int64_t row, int32_t which, char* out, int32_t cap) {
  const std::u16string& s = which == 0 ? t->rows[row]->TaskName : which == 2 ? t->rows[row]->TaskAssignedTo : t->rows[row]->TaskCreatedBy;
  std::string u;
  for (size_t i = 0; i < s.size(); ++i) {
    uint32_t c = s[i];
    if (c >= 0xD800 && c <= 0xDBFF && i + 1 < s.size()) { c = 0x10000 + ((c - 0xD800) << 10) + (s[i + 1] - 0xDC00); ++i; }
    if (c < 0x80) u.push_back((char)c);
    else if (c < 0x800) { u.push_back((char)(0xC0 | (c >> 6))); u.push_back((char)(0x80 | (c & 0x3F))); }
    else if (c < 0x10000) { u.push_back((char)(0xE0 | (c >> 12))); u.push_back((char)(0x80 | ((c >> 6) & 0x3F))); u.push_back((char)(0x80 | (c & 0x3F))); }
    else { u.push_back((char)(0xF0 | (c >> 18))); u.push_back((char)(0x80 | ((c >> 12) & 0x3F))); u.push_back((char)(0x80 | ((c >> 6) & 0x3F))); u.push_back((char)(0x80 | (c & 0x3F))); }
  }
  int32_t n = (int32_t)std::min<size_t>(u.size(), (size_t)cap);
  memcpy(out, u.data(), n);
  return (int32_t)u.size();
}

// Loads n stored documents (document i = text[off[i], off[i+1])) as the reference would
// deserialise them, in order.  A document may be one element of a JSON array cut at its element
// boundaries: one leading '[' or ',' and one trailing ',' or ']' are ignored.  Returns -1, or the
// index of the first document System.Text.Json would reject (nothing is added then).
int64_t orc_add_documents(orc_table* t, const char* text, const uint64_t* off, int64_t n, int32_t local_offset_min) {
  std::vector<TaskModel*> fresh;
  fresh.reserve(n);
  for (int64_t i = 0; i < n; ++i) {
    const char* p = text + off[i];
    const char* end = text + off[i + 1];
    auto ws = [](char c) { return c == ' ' || c == '\t' || c == '\n' || c == '\r'; };
    while (p < end && ws(*p)) ++p;
    if (p < end && (*p == '[' || *p == ',')) ++p;
    while (end > p && ws(end[-1])) --end;
    if (end > p && (end[-1] == ',' || end[-1] == ']')) --end;
    try {
      fresh.push_back(new TaskModel(DeserializeTaskModel(p, end, local_offset_min)));
    } catch (const JsonParseError&) {
      for (TaskModel* m : fresh) delete m;
      return i;
    }
  }
  for (TaskModel* m : fresh) t->Add(m);
  return -1;
}

// One document per listed row, concatenated; off receives n + 1 offsets.  Returns the total length.
int64_t orc_serialize_documents(orc_table* t, const uint32_t* rows, int64_t n, int32_t flavor, int32_t local_offset_min,
                                char* out, int64_t cap, uint64_t* off) {
  std::string s;
  for (int64_t i = 0; i < n; ++i) {
    if (off) off[i] = s.size();
    JsonWriteTaskModel(s, *t->rows[rows[i]], flavor, local_offset_min);
  }
  if (off) off[n] = s.size();
  if (out && cap > 0) memcpy(out, s.data(), (size_t)std::min<int64_t>(cap, (int64_t)s.size()));
  return (int64_t)s.size();
}

void orc_guid_string(const uint8_t* id16, char* out37) {
  Guid g;
  memcpy(g.b, id16, 16);
  std::string s = g.ToString();
  memcpy(out37, s.c_str(), 37);
}

// ---- full-size checker: the same predicates applied to rows generated on the fly ----------
// (no table is stored; used for the 1 B-row parity properties).  Rows are independent, so the
// range is split over threads and the partial sums combined; every quantity is order-free.
typedef struct orc_stream_out {
  int64_t n_listed;           // rows GetYesterdaysDueTasks returns
  int64_t n_marked;           // rows whose IsOverDue the cron tick sets
  uint64_t xor_id_hi, xor_id_lo;  // XOR of the listed TaskIds
  uint64_t sum_rows;          // sum of listed shard-local row numbers (mod 2^64)
  int64_t overdue_after;      // popcount of IsOverDue after the tick
  int64_t completed;          // popcount of IsCompleted
  int32_t threw;              // FAKE mode with survivors
} orc_stream_out;

void orc_stream_sweep(const tt_synth_cfg* cfg, uint64_t first_row, uint64_t n, int32_t mode, int64_t today_local_ticks,
                      int64_t run_at_utc_ticks, int32_t threads, orc_stream_out* out) {
  if (threads < 1) threads = 1;
  std::vector<orc_stream_out> part(threads);
  std::vector<std::thread> pool;
  DateTime yesterday = DateTime(today_local_ticks, TT_KIND_LOCAL).AddDays(-1);
  DateTime runAt(run_at_utc_ticks, TT_KIND_UTC);
  for (int th = 0; th < threads; ++th) {
    pool.emplace_back([&, th]() {
      orc_stream_out o;
      memset(&o, 0, sizeof o);
      uint64_t lo = n * th / threads, hi = n * (th + 1) / threads;
      for (uint64_t i = lo; i < hi; ++i) {
        tt_synth_row r = tt_synth_make(cfg, first_row + i);
        DateTime due(r.due_raw);
        bool listed;
        if (mode == ORC_MODE_FAKE) listed = due.Equals(yesterday);
        else listed = due.Ticks() == yesterday.Ticks() && due.Kind() == TT_KIND_UNSPECIFIED && !r.completed && !r.overdue;
        bool overdue = r.overdue;
        if (listed) {
          o.n_listed++;
          o.xor_id_hi ^= r.id_hi;
          o.xor_id_lo ^= r.id_lo;
          o.sum_rows += i;
          if (runAt.Date() > due.Date()) {
            o.n_marked++;
            if (mode == ORC_MODE_STORE) overdue = true;
          }
        }
        o.overdue_after += overdue;
        o.completed += r.completed;
      }
      part[th] = o;
    });
  }
  for (auto& t : pool) t.join();
  memset(out, 0, sizeof *out);
  for (auto& o : part) {
    out->n_listed += o.n_listed;
    out->n_marked += o.n_marked;
    out->xor_id_hi ^= o.xor_id_hi;
    out->xor_id_lo ^= o.xor_id_lo;
    out->sum_rows += o.sum_rows;
    out->overdue_after += o.overdue_after;
    out->completed += o.completed;
  }
  if (mode == ORC_MODE_FAKE) {
    out->threw = out->n_marked > 0;
    // the throw happens before any state change; nothing is marked in FAKE mode
  }
}

// The creator filter at full size (BASELINE.json config 4: 100 M rows, 1 M creators): for every query the
// rows whose TaskCreatedBy equals it — FakeTasksManager.cs:70 `t.TaskCreatedBy.Equals(createdBy)`, ordinal
// UTF-16 equality (TasksStoreManager.cs:56-59 asks the store for the same EQ) — summarised order-free: count,
// XOR of TaskIds, sum / min / max of the row numbers.  Rows are generated on the fly; each row's creator
// STRING is built and looked up among the query strings (no shortcut through the generator's index).
typedef struct orc_creator_out {
  int64_t count;
  uint64_t xor_id_hi, xor_id_lo;
  uint64_t sum_rows;
  int64_t min_row, max_row;  // -1 when there is no match
} orc_creator_out;

void orc_stream_creators(const tt_synth_cfg* cfg, uint64_t first_row, uint64_t n, const char* pool, const uint32_t* off,
                         int32_t nq, int32_t threads, orc_creator_out* out) {
  if (threads < 1) threads = 1;
  std::unordered_map<std::u16string, std::vector<int32_t>> index;  // a query string may be asked more than once
  for (int32_t q = 0; q < nq; ++q) index[Utf8ToUtf16(pool + off[q], (int32_t)(off[q + 1] - off[q]))].push_back(q);
  std::vector<std::vector<orc_creator_out>> part(threads);
  std::vector<std::thread> workers;
  for (int th = 0; th < threads; ++th) {
    workers.emplace_back([&, th]() {
      std::vector<orc_creator_out> o(nq);
      for (auto& x : o) {
        memset(&x, 0, sizeof x);
        x.min_row = x.max_row = -1;
      }
      const uint64_t lo = n * th / threads, hi = n * (th + 1) / threads;
      char buf[48];
      std::u16string who;
      for (uint64_t i = lo; i < hi; ++i) {
        const tt_synth_row r = tt_synth_make(cfg, first_row + i);
        const int len = tt_synth_creator(r.creator, buf);
        who.assign(buf, buf + len);  // the generator's creator strings are ASCII
        auto it = index.find(who);
        if (it == index.end()) continue;
        for (int32_t q : it->second) {
          orc_creator_out& x = o[q];
          x.count++;
          x.xor_id_hi ^= r.id_hi;
          x.xor_id_lo ^= r.id_lo;
          x.sum_rows += i;
          if (x.min_row < 0) x.min_row = (int64_t)i;
          x.max_row = (int64_t)i;
        }
      }
      part[th] = std::move(o);
    });
  }
  for (auto& w : workers) w.join();
  for (int32_t q = 0; q < nq; ++q) {
    orc_creator_out t;
    memset(&t, 0, sizeof t);
    t.min_row = t.max_row = -1;
    for (int th = 0; th < threads; ++th) {
      const orc_creator_out& x = part[th][q];
      if (x.count == 0) continue;
      t.count += x.count;
      t.xor_id_hi ^= x.xor_id_hi;
      t.xor_id_lo ^= x.xor_id_lo;
      t.sum_rows += x.sum_rows;
      if (t.min_row < 0) t.min_row = x.min_row;  // the threads' ranges ascend
      t.max_row = x.max_row;
    }
    out[q] = t;
  }
}

// ---- "best-effort CPU" context number: the same sweep over SoA columns on all cores -------
// (not the reference's algorithm layout; reported beside the LINQ-faithful number)
typedef struct orc_soa {
  int64_t n;
  uint64_t* due_raw;
  uint8_t* completed;
  uint8_t* overdue;
} orc_soa;

orc_soa* orc_soa_create(const tt_synth_cfg* cfg, uint64_t first_row, int64_t n, int32_t threads) {
  orc_soa* s = new orc_soa();
  s->n = n;
  s->due_raw = new uint64_t[n];
  s->completed = new uint8_t[n];
  s->overdue = new uint8_t[n];
  if (threads < 1) threads = 1;
  std::vector<std::thread> pool;
  for (int th = 0; th < threads; ++th)
    pool.emplace_back([=]() {
      for (int64_t i = n * th / threads; i < n * (th + 1) / threads; ++i) {
        tt_synth_row r = tt_synth_make(cfg, first_row + (uint64_t)i);
        s->due_raw[i] = r.due_raw;
        s->completed[i] = r.completed;
        s->overdue[i] = r.overdue;
      }
    });
  for (auto& t : pool) t.join();
  return s;
}

void orc_soa_reset(orc_soa* s, const tt_synth_cfg* cfg, uint64_t first_row, int32_t threads) {
  if (threads < 1) threads = 1;
  std::vector<std::thread> pool;
  int64_t n = s->n;
  for (int th = 0; th < threads; ++th)
    pool.emplace_back([=]() {
      for (int64_t i = n * th / threads; i < n * (th + 1) / threads; ++i)
        s->overdue[i] = tt_synth_make(cfg, first_row + (uint64_t)i).overdue;
    });
  for (auto& t : pool) t.join();
}

void orc_soa_destroy(orc_soa* s) {
  delete[] s->due_raw;
  delete[] s->completed;
  delete[] s->overdue;
  delete s;
}

// STORE-mode tick over the SoA copy: list + mark.  Returns the hit count; out_rows (cap >= n
// is not required: hits beyond cap are counted but not stored).
int64_t orc_soa_sweep(orc_soa* s, int64_t today_local_ticks, int64_t run_at_utc_ticks, int32_t threads,
                      uint32_t* out_rows, int64_t cap) {
  if (threads < 1) threads = 1;
  const uint64_t yraw = (uint64_t)(today_local_ticks - kTicksPerDay);  // Kind Unspecified = top bits 0
  const bool mark = run_at_utc_ticks / kTicksPerDay > (today_local_ticks - kTicksPerDay) / kTicksPerDay;
  std::atomic<int64_t> cursor{0};
  std::vector<std::thread> pool;
  for (int th = 0; th < threads; ++th)
    pool.emplace_back([&, th]() {
      std::vector<uint32_t> local;
      int64_t lo = s->n * th / threads, hi = s->n * (th + 1) / threads;
      for (int64_t i = lo; i < hi; ++i) {
        if (s->due_raw[i] == yraw && !s->completed[i] && !s->overdue[i]) {
          local.push_back((uint32_t)i);
          if (mark) s->overdue[i] = 1;
        }
      }
      int64_t at = cursor.fetch_add((int64_t)local.size());
      for (size_t k = 0; k < local.size(); ++k)
        if (at + (int64_t)k < cap) out_rows[at + k] = local[k];
    });
  for (auto& t : pool) t.join();
  return cursor.load();
}

}  // extern "C"
