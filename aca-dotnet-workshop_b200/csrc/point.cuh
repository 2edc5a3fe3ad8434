// point.cuh — the single-row operations of ITasksManager (Api/Services/ITasksManager.cs:8-12):
// GetTaskById, CreateNewTask, UpdateTask, MarkTaskCompleted, DeleteTask.  They are not
// data-parallel and are kept deliberately simple; they exist so that the table can stand behind
// the whole interface, not only behind the sweep and the creator filter.
//   FakeTasksManager.cs:32-100 — `_tasksList.FirstOrDefault(t => t.TaskId.Equals(taskId))` then a
//   field update / Remove; TasksStoreManager.cs:22-102 — Get/Save/DeleteStateAsync by key.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include "creator.cuh"
#include "tt_bits.h"

namespace tt {

constexpr uint32_t kRowDeleted = 0xFFFFFFFFu;  // creator_ref of a deleted row

// FirstOrDefault(t => t.TaskId.Equals(id)): the lowest live row with this id, or ~0.
// sorted != 0: rows (or id_order) ascend by TaskId -> one thread does a binary search;
// otherwise every thread scans a slice and the lowest match wins.
__global__ void find_id_kernel(uint4 key, const uint4* task_id, const uint32_t* creator_ref, const uint32_t* id_order,
                               int64_t n, int sorted, unsigned int* out_row) {
  if (sorted) {
    if (blockIdx.x != 0 || threadIdx.x != 0) return;
    int64_t lo = 0, hi = n;
    while (lo < hi) {
      int64_t mid = (lo + hi) >> 1;
      uint32_t r = id_order ? id_order[mid] : (uint32_t)mid;
      if (cmp_id(task_id[r], key) < 0) lo = mid + 1; else hi = mid;
    }
    for (; lo < n; ++lo) {  // skip deleted rows that still carry the id
      uint32_t r = id_order ? id_order[lo] : (uint32_t)lo;
      if (cmp_id(task_id[r], key) != 0) break;
      if (creator_ref[r] != kRowDeleted) {
        *out_row = r;
        break;
      }
    }
    return;
  }
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const uint4 v = task_id[i];
    if (v.x == key.x && v.y == key.y && v.z == key.z && v.w == key.w && creator_ref[i] != kRowDeleted)
      atomicMin(out_row, (unsigned int)i);
  }
}

struct RowPoke {
  uint32_t row;
  uint32_t set_id, set_created, set_due, set_creator, set_completed, set_overdue;
  uint4 id;
  uint64_t created_raw, due_raw, creator_hash;
  uint32_t due_code, creator_ref, completed, overdue;
};

// Writes the selected fields of one row (one thread).
__global__ void poke_row_kernel(RowPoke p, uint32_t* due_code, uint32_t* completed, uint32_t* overdue,
                                uint64_t* creator_hash, uint4* task_id, uint64_t* created_raw, uint64_t* due_raw,
                                uint32_t* creator_ref) {
  const uint32_t r = p.row;
  if (p.set_id) task_id[r] = p.id;
  if (p.set_created) created_raw[r] = p.created_raw;
  if (p.set_due) {
    due_raw[r] = p.due_raw;
    due_code[r] = p.due_code;
  }
  if (p.set_creator) {
    creator_ref[r] = p.creator_ref;
    creator_hash[r] = p.creator_hash;
  }
  const uint32_t bit = 1u << (r & 31);
  if (p.set_completed) {
    if (p.completed) atomicOr(completed + (r >> 5), bit); else atomicAnd(completed + (r >> 5), ~bit);
  }
  if (p.set_overdue) {
    if (p.overdue) atomicOr(overdue + (r >> 5), bit); else atomicAnd(overdue + (r >> 5), ~bit);
  }
}

}  // namespace tt
