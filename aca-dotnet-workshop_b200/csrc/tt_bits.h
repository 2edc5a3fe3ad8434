// tt_bits.h — column encodings and the warp bit-shuffles of the sweep, as plain functions
// usable from host code too (tests/ simulates the 32 lanes on the CPU).
#ifndef TT_BITS_H
#define TT_BITS_H

#include <stdint.h>

#include "../../include/tt_synth.h"

// ---- due_code: the hot 32-bit encoding of TaskDueDate ----------------------
//   bits  0..21  day number  = ticks / TicksPerDay   (max 3 652 058 for year 9999 < 2^22)
//   bit   22     1 if the value carries a time of day (ticks % TicksPerDay != 0)
//   bits 23..24  DateTimeKind (0 Unspecified, 1 Utc, 2 Local)
//   0xFFFFFFFF   padding rows (never match: bits above 24 are set)
// FAKE mode  (FakeTasksManager.cs:109, tick equality)             : (code & TT_MASK_FAKE)  == yesterday
// STORE mode (TasksStoreManager.cs:119-128, JSON string equality) : (code & TT_MASK_STORE) == yesterday
#define TT_DUE_DAY_MASK 0x003FFFFFu
#define TT_DUE_TOD_BIT 0x00400000u
#define TT_DUE_KIND_SHIFT 23
#define TT_MASK_FAKE 0xFE7FFFFFu   /* ignores Kind only */
#define TT_MASK_STORE 0xFFFFFFFFu
#define TT_DUE_PAD 0xFFFFFFFFu

TT_HD uint32_t tt_pack_due(uint64_t raw) {
  uint64_t ticks = raw & TT_TICKS_MASK;
  uint32_t kind = (uint32_t)(raw >> 62);
  uint64_t day = ticks / (uint64_t)TT_TICKS_PER_DAY;
  uint64_t tod = ticks - day * (uint64_t)TT_TICKS_PER_DAY;
  return (uint32_t)day | (tod ? TT_DUE_TOD_BIT : 0u) | (kind << TT_DUE_KIND_SHIFT);
}

// ---- sweep tile geometry ------------------------------------------------------------------
#define TT_TILE_ROWS 8192        /* rows per CTA tile = 256 flag words */
#define TT_TILE_WORDS 256
#define TT_WARP_ROWS 1024        /* rows per warp per tile = 32 flag words, 8 x LDG.128 per lane */

// One butterfly step of the 8x8 nibble transpose across the 8 lanes of a lane group.
// x = own packed nibbles, y = the value of lane (lane ^ d); d in {4, 2, 1}.
TT_HD uint32_t tt_nibble_xpose_step(uint32_t x, uint32_t y, uint32_t lane, uint32_t d) {
  uint32_t m = d == 4 ? 0x0000FFFFu : (d == 2 ? 0x00FF00FFu : 0x0F0F0F0Fu);
  uint32_t sh = 4u * d;
  return (lane & d) ? ((x & ~m) | ((y & ~m) >> sh)) : ((x & m) | ((y & m) << sh));
}

// After the transpose, lane 8g+j holds flag word 4j+g of the warp's 32 words; lane t wants
// word t, which sits in lane tt_word_src_lane(t).
TT_HD uint32_t tt_word_src_lane(uint32_t t) { return 8u * (t & 3u) + (t >> 2); }

#endif
