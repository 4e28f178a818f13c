// Value interning on the device (interner.cu): byte strings -> dense Int32 ids, key bytes in a side arena.
#pragma once
#include <memory>

#include "internal.h"

namespace bgpu {
struct Interner;
int interner_create(bgpu_ctx* ctx, std::shared_ptr<Interner>* out);
uint32_t interner_size(const Interner* in);
// forget every value, keep the capacity
int interner_clear(Interner* in);
// d_ids[r] = id of value r of `col` (utf8 / binary / fixed-width read as bytes), -1 for null rows. Synchronises the stream.
// seed: the interner is empty and the values are known to be pairwise distinct: value r gets id r.
int interner_intern(Interner* in, const Column& col, int64_t n, int32_t* d_ids, bool seed);
// values of ids[0 .. n) as a column shaped like `proto`; ids < 0 or a clear bit in valid_words (32-bit words, may be null) = null
int interner_take(Interner* in, const int32_t* d_ids, const uint32_t* valid_words, int64_t n, const Column& proto, Column* out);
// all interned values in id order
int interner_export(Interner* in, const Column& proto, Column* out);
}  // namespace bgpu
