/* ORACLE (test infrastructure only — never linked into or called by the product path).
 *
 * Timed CPU restatement of the reference's TPC-H Q1 hot path, executed the way Ballista standalone executes it
 * (the Rust reference cannot be built here: no cargo/rustc). One thread per input partition (executor
 * `concurrent_tasks`, /root/reference/ballista/executor/executor_config_spec.toml:75-79), each streaming 8192-row
 * batches (`ballista.batch.size`, /root/reference/ballista/core/src/config.rs:180-182) through, per batch:
 *   FilterExec         l_shipdate <= date: build the selection and materialise the filtered columns
 *   expressions        (1 - l_discount), l_extendedprice * (..), (1 + l_tax), .. * (..) as whole-column Decimal128
 *                      kernels with i128 overflow checks (arrow-arith numeric.rs decimal_op)
 *   group keys         two Utf8 columns -> row bytes -> hash table probe -> dense group index per row
 *   accumulators       one vectorised update per aggregate (sum i128, count)
 * then per task: emit the Partial state, create_hashes(keys; ahash seeds 0,0,0,0) % n -> per-partition state rows
 * (ShuffleWriterExec, /root/reference/ballista/core/src/execution_plans/shuffle_writer.rs:336-391); then one
 * FinalPartitioned merge per output partition (planner.rs:385-408) run on the same thread pool.
 * Also holds the deterministic synthetic lineitem generator shared by the tests (SURVEY.md §8d distributions).
 */
#include <stdint.h>
#include <string.h>

#include <atomic>
#include <chrono>
#include <string>
#include <thread>
#include <unordered_map>
#include <vector>

#include "ahash_fallback.h"
#include "oracle.h"

typedef __int128 i128;

static inline uint64_t splitmix64(uint64_t x) {
  x += 0x9e3779b97f4a7c15ULL;
  x = (x ^ (x >> 30)) * 0xbf58476d1ce4e5b9ULL;
  x = (x ^ (x >> 27)) * 0x94d049bb133111ebULL;
  return x ^ (x >> 31);
}

/* Row r of the synthetic lineitem for `seed`: every value is a pure function of (seed, r). Columns: quantity,
 * extendedprice, discount, tax as Decimal128(15,2) (i128 little endian), returnflag / linestatus as one byte each
 * (Utf8 offsets are 0..n), shipdate as Date32. */
extern "C" int orc_gen_lineitem(int64_t row0, int64_t n, uint64_t seed, void* qty, void* price, void* disc, void* tax,
                                uint8_t* rf, uint8_t* ls, int32_t* ship) {
  i128 *q = (i128*)qty, *p = (i128*)price, *d = (i128*)disc, *t = (i128*)tax;
  for (int64_t i = 0; i < n; i++) {
    const uint64_t base = splitmix64(seed ^ splitmix64((uint64_t)(row0 + i)));
    const uint64_t a = splitmix64(base + 1), b = splitmix64(base + 2), c = splitmix64(base + 3);
    q[i] = (i128)((a % 50 + 1) * 100);
    p[i] = (i128)(90000 + (a >> 20) % (10495000 - 90000 + 1));
    d[i] = (i128)(b % 11);
    t[i] = (i128)((b >> 16) % 9);
    const int32_t sd = 8036 + (int32_t)((b >> 32) % (10561 - 8036 + 1)); /* 1992-01-02 .. 1998-12-01 */
    const int32_t receipt = sd + 1 + (int32_t)(c % 30);
    ship[i] = sd;
    rf[i] = receipt <= 9298 ? ((c >> 8) & 1 ? 'R' : 'A') : 'N'; /* 1995-06-17 */
    ls[i] = sd > 9298 ? 'O' : 'F';
  }
  return 0;
}

namespace {
constexpr int kBatch = 8192;
constexpr int kAggs = 5;   /* sums: qty, price, disc_price, charge, disc */

struct State {              /* Partial state of one task: keys + 5 i128 sums + count */
  std::vector<std::string> keys;
  std::vector<i128> sum[kAggs];
  std::vector<int64_t> count;
  std::unordered_map<std::string, int32_t> index;
  int32_t intern(const std::string& k) {
    auto it = index.find(k);
    if (it != index.end()) return it->second;
    const int32_t id = (int32_t)keys.size();
    index.emplace(k, id);
    keys.push_back(k);
    for (auto& s : sum) s.push_back(0);
    count.push_back(0);
    return id;
  }
};

struct Cols {
  const i128 *qty, *price, *disc, *tax;
  const uint8_t *rf, *ls;
  const int32_t* ship;
};

bool mul_checked(i128 a, i128 b, i128* out) { return !__builtin_mul_overflow(a, b, out); }

/* one input partition: rows [lo, hi) */
int run_task(const Cols& c, int64_t lo, int64_t hi, int32_t date, State* st) {
  std::vector<int32_t> sel(kBatch);
  std::vector<i128> f_qty(kBatch), f_price(kBatch), f_disc(kBatch), f_tax(kBatch), t1(kBatch), disc_price(kBatch), t2(kBatch),
      charge(kBatch);
  std::vector<uint8_t> f_rf(kBatch), f_ls(kBatch);
  std::vector<int32_t> gid(kBatch);
  std::string key;
  for (int64_t b0 = lo; b0 < hi; b0 += kBatch) {
    const int n = (int)std::min<int64_t>(kBatch, hi - b0);
    /* FilterExec: predicate -> selection -> filter_record_batch */
    int m = 0;
    for (int i = 0; i < n; i++)
      if (c.ship[b0 + i] <= date) sel[m++] = i;
    for (int i = 0; i < m; i++) {
      const int64_t r = b0 + sel[i];
      f_qty[i] = c.qty[r]; f_price[i] = c.price[r]; f_disc[i] = c.disc[r]; f_tax[i] = c.tax[r];
      f_rf[i] = c.rf[r]; f_ls[i] = c.ls[r];
    }
    /* expressions, one kernel per operator as arrow-arith does */
    for (int i = 0; i < m; i++) { if (__builtin_sub_overflow((i128)100, f_disc[i], &t1[i])) return 5; }
    for (int i = 0; i < m; i++) { if (!mul_checked(f_price[i], t1[i], &disc_price[i])) return 5; }
    for (int i = 0; i < m; i++) { if (__builtin_add_overflow((i128)100, f_tax[i], &t2[i])) return 5; }
    for (int i = 0; i < m; i++) { if (!mul_checked(disc_price[i], t2[i], &charge[i])) return 5; }
    /* GroupValuesRows: row bytes = (len, bytes) per key column */
    for (int i = 0; i < m; i++) {
      key.assign({(char)1, (char)f_rf[i], (char)1, (char)f_ls[i]});
      gid[i] = st->intern(key);
    }
    /* GroupsAccumulators */
    const i128* vals[kAggs] = {f_qty.data(), f_price.data(), disc_price.data(), charge.data(), f_disc.data()};
    for (int a = 0; a < kAggs; a++) {
      i128* s = st->sum[a].data();
      const i128* v = vals[a];
      for (int i = 0; i < m; i++) s[gid[i]] = (i128)((unsigned __int128)s[gid[i]] + (unsigned __int128)v[i]);
    }
    int64_t* cnt = st->count.data();
    for (int i = 0; i < m; i++) cnt[gid[i]]++;
  }
  return 0;
}

uint64_t key_hash(const std::string& k) { /* create_hashes over the two Utf8 key columns */
  const uint64_t h0 = ah_hash_str((const uint8_t*)&k[1], 1);
  return df_combine_hashes(ah_hash_str((const uint8_t*)&k[3], 1), h0);
}
}  // namespace

/* result rows: up to max_groups; per group: rf, ls, 5 i128 sums (qty, price, disc_price, charge, disc), count.
 * Returns 0, or 5 on arithmetic overflow. */
extern "C" int orc_q1_pipeline(const void* qty, const void* price, const void* disc, const void* tax, const uint8_t* rf,
                               const uint8_t* ls, const int32_t* ship, int64_t rows, int32_t date, int32_t threads,
                               uint32_t partitions, double* seconds, int32_t* n_groups, int32_t max_groups,
                               uint8_t* out_rf, uint8_t* out_ls, void* out_sums /* i128[max_groups][5] */,
                               int64_t* out_count, uint32_t* out_partition) {
  Cols c{(const i128*)qty, (const i128*)price, (const i128*)disc, (const i128*)tax, rf, ls, ship};
  if (threads < 1) threads = 1;
  const auto t0 = std::chrono::steady_clock::now();
  std::vector<State> states((size_t)threads);
  std::vector<int> rc((size_t)threads, 0);
  {
    std::vector<std::thread> pool;
    for (int t = 0; t < threads; t++)
      pool.emplace_back([&, t] {
        const int64_t lo = rows * t / threads, hi = rows * (t + 1) / threads;
        rc[(size_t)t] = run_task(c, lo, hi, date, &states[(size_t)t]);
      });
    for (auto& th : pool) th.join();
  }
  for (int r : rc)
    if (r) return r;
  /* shuffle write: each task's state rows go to partition hash % n; shuffle read + FinalPartitioned per partition */
  std::vector<State> finals((size_t)partitions);
  {
    std::atomic<uint32_t> next{0};
    std::vector<std::thread> pool;
    const int workers = (int)std::min<uint32_t>((uint32_t)threads, partitions);
    for (int w = 0; w < workers; w++)
      pool.emplace_back([&] {
        for (uint32_t p = next++; p < partitions; p = next++) {
          State& f = finals[p];
          for (const State& s : states)
            for (size_t g = 0; g < s.keys.size(); g++) {
              if (key_hash(s.keys[g]) % partitions != p) continue;
              const int32_t id = f.intern(s.keys[g]);
              for (int a = 0; a < kAggs; a++)
                f.sum[a][(size_t)id] = (i128)((unsigned __int128)f.sum[a][(size_t)id] + (unsigned __int128)s.sum[a][g]);
              f.count[(size_t)id] += s.count[g];
            }
        }
      });
    for (auto& th : pool) th.join();
  }
  *seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
  int32_t n = 0;
  for (uint32_t p = 0; p < partitions; p++)
    for (size_t g = 0; g < finals[p].keys.size(); g++) {
      if (n >= max_groups) break;
      out_rf[n] = (uint8_t)finals[p].keys[g][1];
      out_ls[n] = (uint8_t)finals[p].keys[g][3];
      for (int a = 0; a < kAggs; a++) ((i128*)out_sums)[(size_t)n * kAggs + a] = finals[p].sum[a][g];
      out_count[n] = finals[p].count[g];
      out_partition[n] = p;
      n++;
    }
  *n_groups = n;
  return 0;
}
