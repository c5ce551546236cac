/*!
 * rabit-b200 test: the custom-reducer entry points across ranks (what XGBoost actually calls).
 *   Reducer<GradPair, PairSum>     16-byte struct, opaque merge function: tree for small counts, ring for large
 *                                  (reference include/rabit/internal/rabit-inl.h:255-298, ReducerSafe_ path)
 *   Reducer<...>::DeclarePackedSum the same struct declared as a lane-wise double sum
 *   SerializeReducer<Bag>          variable-size objects in fixed slots (rabit-inl.h:300-362)
 * Every result is compared with its closed form; the model is checkpointed every iteration so that mock kills
 * replay cached results of these ops as well.
 * usage: reducer_multi <ndata> [iters=N] [mock=r,v,s,t ...] [rabit_*=...]
 */
#include <algorithm>
#include <cstdio>
#include <vector>

#include "recover_common.h"

namespace {
struct GradPair { double g, h; };
void PairSum(GradPair &d, const GradPair &s) { d.g += s.g; d.h += s.h; }  // NOLINT(*)
struct Mixed { float lo; int hits; double total; };                         // 16 bytes, lanes of different types
void MixedMerge(Mixed &d, const Mixed &s) {                                 // NOLINT(*)
  d.lo = std::min(d.lo, s.lo); d.hits += s.hits; d.total += s.total;
}
struct Tri { float a, b, c; };                                              // 12 bytes: does not divide a power of two
void TriMerge(Tri &d, const Tri &s) { d.a += s.a; d.b = std::max(d.b, s.b); d.c += s.c; }  // NOLINT(*)
/*! \brief a set of small integers that grows when merged: serialised size differs from rank to rank */
struct Bag {
  std::vector<int> items;
  void Load(rabit::Stream &fi) { fi.Read(&items); }         // NOLINT(*)
  void Save(rabit::Stream &fo) const { fo.Write(items); }   // NOLINT(*)
  void Reduce(const Bag &o, size_t) {
    items.insert(items.end(), o.items.begin(), o.items.end());
    std::sort(items.begin(), items.end());
  }
};
}  // namespace

int main(int argc, char *argv[]) {
  if (argc < 2) { std::printf("Usage: reducer_multi <ndata> [name=value ...]\n"); return 0; }
  const size_t n = static_cast<size_t>(atol(argv[1]));
  const int iters = rtest::ArgInt(argc, argv, "iters", 2);
  rabit::Init(argc, argv);
  const int me = rabit::GetRank(), world = rabit::GetWorldSize();
  rtest::VecModel model;
  int it = rabit::LoadCheckPoint(&model);
  if (it == 0) model.Reset(4, 1.0f);
  for (; it < iters; ++it) {
    for (int declared = 0; declared < 2; ++declared) {
      std::vector<GradPair> hp(n);
      for (size_t i = 0; i < n; ++i) { hp[i].g = static_cast<double>((i + me + it) % 17); hp[i].h = 0.5 * me + it; }
      rabit::Reducer<GradPair, PairSum> red;
      if (declared) red.DeclarePackedSum<double>(2);
      red.Allreduce(hp.data(), n);
      for (size_t i = 0; i < n; ++i) {
        double g = 0, h = 0;
        for (int r = 0; r < world; ++r) { g += static_cast<double>((i + r + it) % 17); h += 0.5 * r + it; }
        rabit::utils::Check(hp[i].g == g && hp[i].h == h, "[%d] pair reducer (declared=%d) mismatch at %lu (iter %d)",
                            me, declared, static_cast<unsigned long>(i), it);  // NOLINT(*)
      }
    }
    {
      std::vector<Mixed> mx(n);
      for (size_t i = 0; i < n; ++i) {
        mx[i].lo = static_cast<float>((i * 7 + me * 3) % 11); mx[i].hits = me + 1; mx[i].total = i + 0.25 * me;
      }
      rabit::Reducer<Mixed, MixedMerge> red;
      red.Allreduce(mx.data(), n);
      for (size_t i = 0; i < n; ++i) {
        float lo = 1e30f; int hits = 0; double total = 0;
        for (int r = 0; r < world; ++r) {
          lo = std::min(lo, static_cast<float>((i * 7 + r * 3) % 11)); hits += r + 1; total += i + 0.25 * r;
        }
        rabit::utils::Check(mx[i].lo == lo && mx[i].hits == hits && mx[i].total == total,
                            "[%d] mixed-lane reducer mismatch at %lu (iter %d)", me, static_cast<unsigned long>(i), it);  // NOLINT(*)
      }
    }
    {
      // items of 12 bytes straddle every power-of-two boundary of a staging or shared-memory ring
      std::vector<Tri> tr(n);
      for (size_t i = 0; i < n; ++i) {
        tr[i].a = static_cast<float>((i + me) % 13); tr[i].b = static_cast<float>((i * (me + 1)) % 29); tr[i].c = 1.0f;
      }
      rabit::Reducer<Tri, TriMerge> red;
      red.Allreduce(tr.data(), n);
      for (size_t i = 0; i < n; ++i) {
        float a = 0, b = 0;
        for (int r = 0; r < world; ++r) { a += static_cast<float>((i + r) % 13); b = std::max(b, static_cast<float>((i * (r + 1)) % 29)); }
        rabit::utils::Check(tr[i].a == a && tr[i].b == b && tr[i].c == static_cast<float>(world),
                            "[%d] 12-byte reducer mismatch at %lu (iter %d)", me, static_cast<unsigned long>(i), it);  // NOLINT(*)
      }
    }
    {
      const size_t nobj = 5;
      std::vector<Bag> bags(nobj);
      for (size_t j = 0; j < nobj; ++j) {
        for (int k = 0; k <= me % 3; ++k) bags[j].items.push_back(static_cast<int>(j * 100 + me * 4 + k));
      }
      rabit::SerializeReducer<Bag> sred;
      sred.Allreduce(bags.data(), 8 + 4 * 3 * static_cast<size_t>(world) + 16, nobj);
      for (size_t j = 0; j < nobj; ++j) {
        std::vector<int> want;
        for (int r = 0; r < world; ++r) {
          for (int k = 0; k <= r % 3; ++k) want.push_back(static_cast<int>(j * 100 + r * 4 + k));
        }
        std::sort(want.begin(), want.end());
        rabit::utils::Check(bags[j].items == want, "[%d] serialize reducer mismatch in object %lu (iter %d)", me,
                            static_cast<unsigned long>(j), it);  // NOLINT(*)
      }
    }
    rabit::CheckPoint(&model);
    std::printf("[%d] iter %d checkpointed (version %d)\n", me, it, rabit::VersionNumber());
  }
  rabit::Finalize();
  return 0;
}
