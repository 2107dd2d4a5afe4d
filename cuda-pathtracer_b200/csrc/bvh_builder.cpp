// bvh_builder.cpp -- host builder for the 8-wide compressed BVH (bvh.h).
//
// Pipeline: padded triangle boxes -> binary BVH by binned SAH (leaves <= 3
// triangles) -> optimal collapse to 8-wide nodes (dynamic program) -> octant-ordered
// slot assignment -> breadth-first emission (internal children contiguous, a node's
// triangles contiguous) with outward-rounded 8-bit quantisation.
//
// New work relative to the reference, which tests every face per ray
// (cuda_opengl/include/shaders/intersection.cuh:179-196).  The reference-side step this
// build replaces at load time is Scene::upload_meshes (cuda_opengl/src/scene/scene.cpp:202-283).
#include "bvh.h"

#include <algorithm>
#include <atomic>
#include <cfloat>
#include <chrono>
#include <cmath>
#include <condition_variable>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <functional>
#include <memory>
#include <mutex>
#include <numeric>
#include <thread>
#include <sys/mman.h>

#if defined(__SSE2__) && !defined(PTC_BUILD_SCALAR) // (aarch64 hosts and -DPTC_BUILD_SCALAR take the plain C++ forms: same bits)
#include <emmintrin.h>
#define PTC_BUILD_SSE 1
#else
#define PTC_BUILD_SSE 0
#endif

namespace ptc {
namespace {

struct Box
{
  float lo[3] = { FLT_MAX, FLT_MAX, FLT_MAX };
  float hi[3] = { -FLT_MAX, -FLT_MAX, -FLT_MAX };
  void grow(const Box& b)
  {
    for (int a = 0; a < 3; ++a) {
      lo[a] = std::min(lo[a], b.lo[a]);
      hi[a] = std::max(hi[a], b.hi[a]);
    }
  }
  void grow(const float* p)
  {
    for (int a = 0; a < 3; ++a) {
      lo[a] = std::min(lo[a], p[a]);
      hi[a] = std::max(hi[a], p[a]);
    }
  }
  float half_area() const
  {
    float dx = hi[0] - lo[0], dy = hi[1] - lo[1], dz = hi[2] - lo[2];
    return dx * dy + dy * dz + dz * dx;
  }
  bool valid() const { return lo[0] <= hi[0]; }
};

struct BinNode
{
  Box box;
  int left = -1, right = -1; // children (internal) ...
  uint32_t first = 0, count = 0; // ... or a range of Builder::ids (leaf)
  bool leaf() const { return left < 0; }
};

constexpr int kBins = 16;
constexpr uint32_t kMaxLeaf = 3;

static float
trav_cost()
{
  static float c = getenv("PTC_TRAV_COST") ? (float)atof(getenv("PTC_TRAV_COST")) : 1.0f;
  return c;
}

// Worker threads for the host build: hardware threads, capped; PTC_BUILD_THREADS overrides.
static unsigned
build_threads()
{
  if (const char* e = getenv("PTC_BUILD_THREADS"))
    return (unsigned)std::max(1, atoi(e));
  const unsigned hw = std::thread::hardware_concurrency();
  return std::min(32u, std::max(1u, hw));
}

// f(lo, hi) over [0, n) in contiguous chunks, one per worker thread (serial when small).
template <typename F>
static void
parallel_for(uint32_t n, F f)
{
  const unsigned t = n >= 4096 ? build_threads() : 1u;
  if (t <= 1) {
    f(0u, n);
    return;
  }
  std::vector<std::thread> th;
  const uint32_t per = (n + t - 1) / t;
  for (unsigned k = 1; k < t; ++k) {
    const uint32_t lo = std::min(n, k * per), hi = std::min(n, (k + 1) * per);
    if (lo < hi)
      th.emplace_back([=, &f] { f(lo, hi); });
  }
  f(0u, std::min(n, per));
  for (std::thread& x : th)
    x.join();
}

// The workers of ONE build: created once, woken per pass (a build of 10 M triangles runs a few
// hundred data-parallel passes; a thread per pass and chunk was measurable).  run(f) calls f(k)
// on every thread k (the caller is thread 0) and returns when all are back.  Not re-entrant.
class Pool
{
public:
  explicit Pool(unsigned n) : n_(std::max(1u, n))
  {
    for (unsigned k = 1; k < n_; ++k)
      th_.emplace_back([this, k] { worker(k); });
  }
  ~Pool()
  {
    {
      std::lock_guard<std::mutex> l(m_);
      stop_ = true;
      ++gen_;
    }
    start_.notify_all();
    for (std::thread& t : th_)
      t.join();
  }
  unsigned size() const { return n_; }

  template <typename F>
  void run(F&& f)
  {
    if (n_ == 1) {
      f(0u);
      return;
    }
    const std::function<void(unsigned)> job = [&f](unsigned k) { f(k); };
    {
      std::lock_guard<std::mutex> l(m_);
      job_ = &job;
      pending_ = n_ - 1;
      ++gen_;
    }
    start_.notify_all();
    f(0u);
    std::unique_lock<std::mutex> l(m_);
    done_.wait(l, [&] { return pending_ == 0; });
    job_ = nullptr;
  }

  // f(lo, hi, k) over [first, last) cut into size() contiguous chunks (chunk k on thread k;
  // short ranges run on the caller as chunk 0)
  template <typename F>
  void chunks(uint32_t first, uint32_t last, F&& f)
  {
    const uint32_t n = last - first;
    if (n_ == 1 || n < 4096) {
      if (n)
        f(first, last, 0u);
      return;
    }
    const uint32_t per = (n + n_ - 1) / n_;
    run([&](unsigned k) {
      const uint32_t lo = first + std::min(n, k * per), hi = first + std::min(n, (k + 1) * per);
      if (lo < hi)
        f(lo, hi, k);
    });
  }

  // f(i) for i in [0, n), handed out one at a time (uneven items: subtrees)
  template <typename F>
  void dynamic(uint32_t n, F&& f)
  {
    std::atomic<uint32_t> next{ 0 };
    run([&](unsigned) {
      for (uint32_t i; (i = next.fetch_add(1u, std::memory_order_relaxed)) < n;)
        f(i);
    });
  }

private:
  void worker(unsigned k)
  {
    uint64_t seen = 0;
    for (;;) {
      const std::function<void(unsigned)>* job;
      {
        std::unique_lock<std::mutex> l(m_);
        start_.wait(l, [&] { return gen_ != seen; });
        seen = gen_;
        if (stop_)
          return;
        job = job_;
      }
      (*job)(k);
      {
        std::lock_guard<std::mutex> l(m_);
        if (--pending_ == 0)
          done_.notify_one();
      }
    }
  }
  const unsigned n_;
  std::vector<std::thread> th_;
  std::mutex m_;
  std::condition_variable start_, done_;
  const std::function<void(unsigned)>* job_ = nullptr;
  uint64_t gen_ = 0;
  unsigned pending_ = 0;
  bool stop_ = false;
};

// malloc'ed array that is NOT value-initialised: the thread that fills a range is the one that
// first touches its pages (one thread zero-filling 0.4 - 1 GB arrays was a third of the old build).
// Big arrays ask for transparent huge pages (where the kernel grants them on madvise): the collapse
// and the emission walk them at random.
template <typename T>
struct Raw
{
  T* p = nullptr;
  Raw() = default;
  Raw(const Raw&) = delete;
  Raw& operator=(const Raw&) = delete;
  ~Raw() { std::free(p); }
  void alloc(size_t count)
  {
    std::free(p);
    const size_t bytes = std::max<size_t>(1, count) * sizeof(T);
    if (bytes >= (64u << 20)) {
      void* q = nullptr;
      if (posix_memalign(&q, 2u << 20, (bytes + (2u << 20) - 1) & ~(size_t)((2u << 20) - 1)) != 0)
        throw std::bad_alloc();
      madvise(q, bytes, MADV_HUGEPAGE);
      p = static_cast<T*>(q);
    } else
      p = static_cast<T*>(std::malloc(bytes));
    if (!p)
      throw std::bad_alloc();
  }
  void release()
  {
    std::free(p);
    p = nullptr;
  }
  T& operator[](size_t i) { return p[i]; }
  const T& operator[](size_t i) const { return p[i]; }
};

// Binned-SAH binary tree, built top-down (scene.cpp:202-283 builds nothing, so at 10 M triangles
// this step IS the load time).  Two phases:
//  1. "wide" nodes (more than `wide_min` primitives) are handled one at a time by ALL threads:
//     every thread bins its slice of the node's range, and a stable two-buffer scatter partitions
//     it (the per-thread bin counts of the chosen axis already say how many primitives of each
//     slice go left, so no counting pass is needed);
//  2. the subtrees below them are independent tasks, handed to the threads largest first; each is
//     the plain serial recursion, in place, with node slots from a range reserved for the task
//     (no shared counter: 20 M fetch_adds on one line serialise 16 threads).
// A node costs TWO passes over its primitives: the bins of the chosen axis also hold the boxes of
// the two children (min / max are exact, so this is the box a separate pass would compute), and
// the children's centroid boxes are gathered while partitioning.
// The tree does not depend on the thread count or on timing: bins, boxes and counts are
// min / max / integer sums, a split depends only on WHICH primitives a node holds, and the one
// place where their order would matter (coincident centroids: a positional median split) orders
// them by face id first.  Leaves list their faces in ascending id (emission).
struct Builder
{
  struct Prim
  {
    Box box;
    float c[3]; // centroid
    uint32_t id;
  };
  static_assert(sizeof(Prim) == 40, "Prim is read as two overlapping 16-byte loads");

  // bins of one node (or of one thread's slice of it), all three axes filled in one pass
  struct Bins
  {
#if PTC_BUILD_SSE
    __m128 lo[3][kBins], hi[3][kBins]; // lanes 0..2 = x, y, z; lane 3 is scratch
#else
    Box box[3][kBins];
#endif
    uint32_t cnt[3][kBins];
    uint32_t occupied[3]; // bit b: bin b of the axis holds something
    Bins()
    {
      std::memset(cnt, 0, sizeof(cnt));
      occupied[0] = occupied[1] = occupied[2] = 0u;
#if PTC_BUILD_SSE
      const __m128 big = _mm_set1_ps(FLT_MAX), small = _mm_set1_ps(-FLT_MAX);
      for (int a = 0; a < 3; ++a)
        for (int b = 0; b < kBins; ++b)
          lo[a][b] = big, hi[a][b] = small;
#endif
    }
    // operand order: min_ps(x, acc) / max_ps(x, acc) return acc unless x is strictly smaller /
    // larger -- exactly std::min(acc, x) / std::max(acc, x), signed zeros included
    void add(int a, int b, const Prim& p)
    {
#if PTC_BUILD_SSE
      lo[a][b] = _mm_min_ps(_mm_loadu_ps(p.box.lo), lo[a][b]);
      hi[a][b] = _mm_max_ps(_mm_loadu_ps(p.box.hi), hi[a][b]); // reads hi[0..2] and c[0]
#else
      box[a][b].grow(p.box);
#endif
      cnt[a][b]++;
      occupied[a] |= 1u << b;
    }
    void merge(const Bins& o)
    {
      for (int a = 0; a < 3; ++a) {
        occupied[a] |= o.occupied[a];
        for (int b = 0; b < kBins; ++b) {
#if PTC_BUILD_SSE
          lo[a][b] = _mm_min_ps(o.lo[a][b], lo[a][b]);
          hi[a][b] = _mm_max_ps(o.hi[a][b], hi[a][b]);
#else
          box[a][b].grow(o.box[a][b]);
#endif
          cnt[a][b] += o.cnt[a][b];
        }
      }
    }
    Box get(int a, int b) const
    {
      Box r;
#if PTC_BUILD_SSE
      alignas(16) float l[4], h[4];
      _mm_store_ps(l, lo[a][b]);
      _mm_store_ps(h, hi[a][b]);
      for (int k = 0; k < 3; ++k)
        r.lo[k] = l[k], r.hi[k] = h[k];
#else
      r = box[a][b];
#endif
      return r;
    }
  };

  // running union of bins for the plane sweep; area() is Box::half_area() operation for operation
  // (the products summed left to right), so both forms give the same bits
  struct Sweep
  {
#if PTC_BUILD_SSE
    __m128 lo = _mm_set1_ps(FLT_MAX), hi = _mm_set1_ps(-FLT_MAX);
    void grow(const Bins& B, int a, int b)
    {
      lo = _mm_min_ps(B.lo[a][b], lo);
      hi = _mm_max_ps(B.hi[a][b], hi);
    }
    float area() const
    {
      const __m128 d = _mm_sub_ps(hi, lo);                               // dx dy dz ..
      const __m128 m = _mm_mul_ps(d, _mm_shuffle_ps(d, d, _MM_SHUFFLE(3, 0, 2, 1))); // dx*dy dy*dz dz*dx ..
      __m128 s = _mm_add_ss(m, _mm_shuffle_ps(m, m, _MM_SHUFFLE(1, 1, 1, 1)));
      s = _mm_add_ss(s, _mm_shuffle_ps(m, m, _MM_SHUFFLE(2, 2, 2, 2)));
      return _mm_cvtss_f32(s);
    }
#else
    Box acc;
    void grow(const Bins& B, int a, int b) { acc.grow(B.box[a][b]); }
    float area() const { return acc.half_area(); }
#endif
  };

  // box of the centroids of the primitives one side of a partition receives
  struct CentroidBox
  {
#if PTC_BUILD_SSE
    __m128 lo = _mm_set1_ps(FLT_MAX), hi = _mm_set1_ps(-FLT_MAX);
    void grow(const Prim& q)
    {
      const __m128 c = _mm_loadu_ps(q.c); // c[0..2] and the id bits (lane 3, ignored)
      lo = _mm_min_ps(c, lo);
      hi = _mm_max_ps(c, hi);
    }
    Box get() const
    {
      alignas(16) float l[4], h[4];
      _mm_store_ps(l, lo);
      _mm_store_ps(h, hi);
      Box r;
      for (int k = 0; k < 3; ++k)
        r.lo[k] = l[k], r.hi[k] = h[k];
      return r;
    }
#else
    Box acc;
    void grow(const Prim& q) { acc.grow(q.c); }
    Box get() const { return acc; }
#endif
  };

  struct Split
  {
    int axis = -1, bin = -1; // primitives whose bin on `axis` is <= `bin` go left
    float cost = FLT_MAX;
    float scale[3], lo[3];
    bool on[3];
    int bin_of(const Prim& p, int a) const { return std::min(kBins - 1, (int)((p.c[a] - lo[a]) * scale[a])); }
  };

  // a subtree that one thread builds (phase 2), or a wide node waiting for phase 1
  struct Job
  {
    uint32_t first, count, depth;
    int src;      // which of the two primitive buffers holds [first, first + count)
    Box box, cbox;
    int parent;   // node that gets this subtree as its left (side 0) or right child; -1 = root
    int side;
    uint32_t node_base; // first of the 2 * count - 1 node slots reserved for the task
  };

  Pool& pool;
  const uint32_t n;
  Raw<Prim> buf[2];
  Raw<BinNode> nodes;      // slots: wide nodes first, then one range per task (gaps stay untouched)
  size_t n_slots = 0;
  Raw<uint32_t> ids;       // position -> face id, what the leaves' [first, first + count) index
  std::vector<Job> tasks;  // the subtrees of phase 2; tasks[i]'s root is node tasks[i].node_base
  uint32_t n_wide = 0;     // nodes [0, n_wide) are the wide nodes, parents before children
  int root = 0;
  uint32_t max_depth = 0;

  Builder(Pool& p, uint32_t n_prims) : pool(p), n(n_prims)
  {
    buf[0].alloc(n);
    ids.alloc(n);
  }

  static void prepare_split(Split& s, const Box& cbox)
  {
    for (int a = 0; a < 3; ++a) {
      const float ext = cbox.hi[a] - cbox.lo[a];
      s.on[a] = ext > 0.f;
      s.scale[a] = s.on[a] ? kBins / ext : 0.f;
      s.lo[a] = cbox.lo[a];
    }
  }

  static void fill_bins(Bins& B, const Split& s, const Prim* p, uint32_t lo, uint32_t hi)
  {
#if PTC_BUILD_SSE
    if (s.on[0] && s.on[1] && s.on[2]) {
      // the three bin indices at once: the same subtract, multiply and truncation per lane as
      // bin_of() (lane 3 carries the id bits through and is ignored); indices are 0..16, so a
      // 16-bit min clamps them
      const __m128 slo = _mm_setr_ps(s.lo[0], s.lo[1], s.lo[2], 0.f);
      const __m128 ssc = _mm_setr_ps(s.scale[0], s.scale[1], s.scale[2], 0.f);
      const __m128i cap = _mm_set1_epi32(kBins - 1);
      for (uint32_t i = lo; i < hi; ++i) {
        const __m128 t = _mm_mul_ps(_mm_sub_ps(_mm_loadu_ps(p[i].c), slo), ssc);
        const __m128i b = _mm_min_epi16(_mm_cvttps_epi32(t), cap);
        const __m128 plo = _mm_loadu_ps(p[i].box.lo), phi = _mm_loadu_ps(p[i].box.hi);
        const int b0 = _mm_cvtsi128_si32(b), b1 = _mm_extract_epi16(b, 2), b2 = _mm_extract_epi16(b, 4);
        B.lo[0][b0] = _mm_min_ps(plo, B.lo[0][b0]);
        B.hi[0][b0] = _mm_max_ps(phi, B.hi[0][b0]);
        B.lo[1][b1] = _mm_min_ps(plo, B.lo[1][b1]);
        B.hi[1][b1] = _mm_max_ps(phi, B.hi[1][b1]);
        B.lo[2][b2] = _mm_min_ps(plo, B.lo[2][b2]);
        B.hi[2][b2] = _mm_max_ps(phi, B.hi[2][b2]);
        B.cnt[0][b0]++, B.cnt[1][b1]++, B.cnt[2][b2]++;
        B.occupied[0] |= 1u << b0, B.occupied[1] |= 1u << b1, B.occupied[2] |= 1u << b2;
      }
      return;
    }
#endif
    for (uint32_t i = lo; i < hi; ++i)
      for (int a = 0; a < 3; ++a)
        if (s.on[a])
          B.add(a, s.bin_of(p[i], a), p[i]);
  }

  // the cheapest of the 3 x 15 candidate planes (first one wins ties: axis, then bin, ascending).
  // Only planes right behind an OCCUPIED bin are evaluated: between two occupied bins the cost does
  // not change and the first plane of such a run is the one a sweep over all 15 would keep (strict
  // <), so the choice is the same -- and most nodes are small, with two to four occupied bins.
  static void choose(Split& s, const Bins& B)
  {
    for (int a = 0; a < 3; ++a) {
      if (!s.on[a])
        continue;
      int occ[kBins], m = 0;
      for (uint32_t bits = B.occupied[a]; bits; bits &= bits - 1)
        occ[m++] = __builtin_ctz(bits);
      float right_area[kBins];
      uint32_t right_cnt[kBins];
      Sweep right;
      uint32_t cnt = 0;
      for (int i = m - 1; i > 0; --i) {
        right.grow(B, a, occ[i]);
        cnt += B.cnt[a][occ[i]];
        right_area[i] = right.area();
        right_cnt[i] = cnt;
      }
      Sweep left;
      cnt = 0;
      for (int i = 0; i + 1 < m; ++i) {
        left.grow(B, a, occ[i]);
        cnt += B.cnt[a][occ[i]];
        const float cost = left.area() * cnt + right_area[i + 1] * right_cnt[i + 1];
        if (cost < s.cost) {
          s.cost = cost;
          s.axis = a;
          s.bin = occ[i];
        }
      }
    }
  }

  static void child_boxes(const Split& s, const Bins& B, Box& left, Box& right)
  {
    for (int b = 0; b < kBins; ++b)
      if (B.cnt[s.axis][b])
        (b <= s.bin ? left : right).grow(B.get(s.axis, b));
  }

  static bool keep_leaf(const Box& box, uint32_t count, const Split& s)
  {
    return count <= kMaxLeaf && box.half_area() * count <= s.cost + box.half_area() * trav_cost();
  }

  // Coincident centroids (no plane separates anything): ordered by face id, cut in the middle.
  static uint32_t median_by_id(Prim* p, uint32_t first, uint32_t count, Box lr[2], Box clr[2])
  {
    std::sort(p + first, p + first + count, [](const Prim& x, const Prim& y) { return x.id < y.id; });
    const uint32_t mid = first + count / 2;
    for (uint32_t i = first; i < first + count; ++i) {
      lr[i >= mid].grow(p[i].box);
      clr[i >= mid].grow(p[i].c);
    }
    return mid;
  }

  // ---- phase 2: one subtree on one thread, in place in buffer p ----------------------------
  struct Local
  {
    uint32_t next_node; // next free slot of the task's range
    uint32_t max_depth = 0;
  };

  // bins + partition of one node; false = it stays a leaf.  Not inlined: its 1.7 KB of bins must
  // not sit in every frame of the recursion below.
  __attribute__((noinline)) static bool split_serial(Prim* p, uint32_t first, uint32_t count, const Box& box,
                                                     const Box& cbox, uint32_t& mid, Box lr[2], Box clr[2])
  {
    Split s;
    prepare_split(s, cbox);
    if (!(s.on[0] || s.on[1] || s.on[2])) {
      if (count <= kMaxLeaf)
        return false;
      mid = median_by_id(p, first, count, lr, clr);
      return true;
    }
    Bins B;
    fill_bins(B, s, p, first, first + count);
    choose(s, B);
    if (keep_leaf(box, count, s))
      return false; // cheaper (or as cheap) to keep as a leaf
    child_boxes(s, B, lr[0], lr[1]);
    // in-place partition (the exchange order of std::partition); every primitive is classified
    // exactly once, and its centroid joins the box of the side it lands on
    const int a = s.axis;
    const float slo = s.lo[a], ssc = s.scale[a];
    const int sbin = s.bin;
    auto goes_left = [=](const Prim& q) { return std::min(kBins - 1, (int)((q.c[a] - slo) * ssc)) <= sbin; };
    CentroidBox cl, cr;
    uint32_t i = first, j = first + count;
    for (;;) {
      while (i < j && goes_left(p[i])) {
        cl.grow(p[i]);
        ++i;
      }
      if (i == j)
        break;
      cr.grow(p[i]);
      --j;
      while (i < j && !goes_left(p[j])) {
        cr.grow(p[j]);
        --j;
      }
      if (i == j)
        break;
      cl.grow(p[j]);
      std::swap(p[i], p[j]);
      ++i;
    }
    clr[0] = cl.get(), clr[1] = cr.get();
    mid = i;
    if (mid == first || mid == first + count) { // cannot happen (both sides of the plane hold a bin)
      lr[0] = lr[1] = clr[0] = clr[1] = Box();
      mid = median_by_id(p, first, count, lr, clr);
    }
    return true;
  }

  int build_serial(Prim* p, uint32_t first, uint32_t count, uint32_t depth, const Box& box, const Box& cbox,
                   Local& L)
  {
    L.max_depth = std::max(L.max_depth, depth);
    const int idx = (int)L.next_node++;
    BinNode* nd = new (&nodes[idx]) BinNode();
    nd->box = box;
    nd->first = first;
    nd->count = count;
    if (count <= 1)
      return idx;
    Box lr[2], clr[2];
    uint32_t mid;
    if (!split_serial(p, first, count, box, cbox, mid, lr, clr))
      return idx;
    const int l = build_serial(p, first, mid - first, depth + 1, lr[0], clr[0], L);
    const int r = build_serial(p, mid, first + count - mid, depth + 1, lr[1], clr[1], L);
    nodes[idx].left = l;
    nodes[idx].right = r;
    return idx;
  }

  // ---- phase 1 + driver -------------------------------------------------------------------
  // `box`, `cbox`: bounds of all primitives / of their centroids (buf[0] is filled by the caller)
  void build(const Box& box, const Box& cbox)
  {
    const unsigned T = pool.size();
    // wide nodes: enough of them that the tasks below are many (8 per thread) and small
    const uint32_t wide_min = T > 1 ? std::max<uint32_t>(8192u, n / (8u * T)) : 0xFFFFFFFFu;
    std::vector<BinNode> top; // the wide nodes, in the order they are taken up
    std::vector<Job> open;    // wide nodes still to split (LIFO)
    Job first_job{ 0u, n, 0u, 0, box, cbox, -1, 0, 0u };
    (n > wide_min ? open : tasks).push_back(first_job);
    if (!open.empty())
      buf[1].alloc(n);
    std::vector<Bins> part(T);
    std::vector<Box> cl(T), cr(T);
    std::vector<uint32_t> n_left(T);
    while (!open.empty()) {
      const Job j = open.back();
      open.pop_back();
      const int idx = (int)top.size();
      top.emplace_back();
      top[idx].box = j.box;
      top[idx].first = j.first;
      top[idx].count = j.count;
      if (j.parent >= 0)
        (j.side ? top[j.parent].right : top[j.parent].left) = idx;
      max_depth = std::max(max_depth, j.depth);

      Split s;
      prepare_split(s, j.cbox);
      Prim* const src = buf[j.src].p;
      Box lr[2], clr[2];
      uint32_t mid;
      int child_src = j.src;
      if (!(s.on[0] || s.on[1] || s.on[2])) {
        mid = median_by_id(src, j.first, j.count, lr, clr); // (a wide node is never a leaf)
      } else {
        for (unsigned k = 0; k < T; ++k)
          part[k] = Bins();
        pool.chunks(j.first, j.first + j.count, [&](uint32_t lo, uint32_t hi, unsigned k) { fill_bins(part[k], s, src, lo, hi); });
        Bins B = part[0];
        for (unsigned k = 1; k < T; ++k)
          B.merge(part[k]);
        choose(s, B);
        child_boxes(s, B, lr[0], lr[1]);
        // stable scatter into the other buffer: slice k writes its left primitives behind those of
        // the slices before it, its right primitives likewise behind `mid`
        uint32_t total_left = 0;
        for (unsigned k = 0; k < T; ++k) {
          n_left[k] = 0;
          for (int b = 0; b <= s.bin; ++b)
            n_left[k] += part[k].cnt[s.axis][b];
          total_left += n_left[k];
          cl[k] = cr[k] = Box();
        }
        mid = j.first + total_left;
        Prim* const dst = buf[1 - j.src].p;
        const int a = s.axis;
        pool.chunks(j.first, j.first + j.count, [&](uint32_t lo, uint32_t hi, unsigned k) {
          uint32_t wl = j.first, wr = mid + (lo - j.first);
          for (unsigned q = 0; q < k; ++q)
            wl += n_left[q];
          wr -= wl - j.first; // right primitives before this slice = primitives before it - left ones
          CentroidBox bl, br;
          for (uint32_t i = lo; i < hi; ++i) {
            if (s.bin_of(src[i], a) <= s.bin) {
              bl.grow(src[i]);
              dst[wl++] = src[i];
            } else {
              br.grow(src[i]);
              dst[wr++] = src[i];
            }
          }
          cl[k] = bl.get(), cr[k] = br.get();
        });
        for (unsigned k = 0; k < T; ++k) {
          clr[0].grow(cl[k]);
          clr[1].grow(cr[k]);
        }
        child_src = 1 - j.src;
      }
      const Job kids[2] = { Job{ j.first, mid - j.first, j.depth + 1, child_src, lr[0], clr[0], idx, 0, 0u },
                            Job{ mid, j.first + j.count - mid, j.depth + 1, child_src, lr[1], clr[1], idx, 1, 0u } };
      for (int c = 1; c >= 0; --c) // (left child taken up first)
        (kids[c].count > wide_min ? open : tasks).push_back(kids[c]);
    }

    // node slots: the wide nodes, then 2 * count - 1 per task
    n_wide = (uint32_t)top.size();
    size_t at = n_wide;
    for (Job& t : tasks) {
      t.node_base = (uint32_t)at;
      at += 2 * (size_t)t.count - 1;
    }
    n_slots = at;
    nodes.alloc(n_slots);
    for (uint32_t i = 0; i < n_wide; ++i)
      new (&nodes[i]) BinNode(top[i]);
    for (const Job& t : tasks)
      if (t.parent >= 0)
        (t.side ? nodes[t.parent].right : nodes[t.parent].left) = (int)t.node_base;
    root = 0; // the first wide node, or the root of the only task

    // phase 2, largest subtree first
    std::vector<uint32_t> order(tasks.size());
    std::iota(order.begin(), order.end(), 0u);
    std::stable_sort(order.begin(), order.end(), [&](uint32_t x, uint32_t y) { return tasks[x].count > tasks[y].count; });
    std::vector<uint32_t> depth_of(tasks.size(), 0u);
    pool.dynamic((uint32_t)tasks.size(), [&](uint32_t q) {
      const Job& t = tasks[order[q]];
      Local L;
      L.next_node = t.node_base;
      Prim* const p = buf[t.src].p;
      build_serial(p, t.first, t.count, t.depth, t.box, t.cbox, L);
      for (uint32_t i = t.first; i < t.first + t.count; ++i)
        ids[i] = p[i].id;
      depth_of[order[q]] = L.max_depth;
    });
    for (uint32_t d : depth_of)
      max_depth = std::max(max_depth, d);
    buf[0].release();
    buf[1].release();
  }
};

// Optimal collapse of the binary tree into 8-wide nodes (dynamic program over
// "represent the subtree of n as a forest of at most i wide-node children", after
// Ylitie, Karras, Laine 2017, section 3.1).  Cost = surface-area-weighted node visits
// (c_node) and triangle tests (c_prim); the ratio follows the measured instruction cost
// of one 8-wide node test (~260) against one triangle test (~55) in the traversal kernel.
struct CollapseDp
{
  static constexpr float kNodeCost = 1.0f;
  const float kPrimCost = getenv("PTC_PRIM_COST") ? (float)atof(getenv("PTC_PRIM_COST")) : 0.22f;
  const BinNode* const nodes;
  struct Entry
  {
    float c[8];        // c[i], i = 1..7
    uint8_t split[9];  // distribute(j): roots given to the left child, j = 2..8
    uint8_t use_dist[8]; // c[i] came from distribute(i) (else from c[i-1]), i = 2..7
    bool leaf1;        // c[1] is the leaf alternative
  };
  Raw<Entry> e; // one per node slot; the entries of used slots are written by solve()
  double sah_sum = 0; // sum over the binary tree of (leaf ? count : 1) * half_area (diagnostic)

  // An entry depends on its two children only: the builder's task subtrees are solved
  // concurrently (bottom-up inside each), then the few wide nodes above them, children first.
  CollapseDp(const Builder& b, Pool& pool) : nodes(b.nodes.p)
  {
    e.alloc(b.n_slots);
    std::vector<uint32_t> order(b.tasks.size());
    std::iota(order.begin(), order.end(), 0u);
    std::stable_sort(order.begin(), order.end(), [&](uint32_t x, uint32_t y) { return b.tasks[x].count > b.tasks[y].count; });
    std::vector<double> part(b.tasks.size(), 0.0);
    pool.dynamic((uint32_t)b.tasks.size(), [&](uint32_t q) {
      double s = 0;
      solve((int)b.tasks[order[q]].node_base, s);
      part[order[q]] = s;
    });
    for (double s : part) // task order, not completion order: the same sum every time
      sah_sum += s;
    for (uint32_t i = b.n_wide; i-- > 0;) // a wide node's children have larger indices
      solve_node((int)i, sah_sum);
  }

  void solve(int idx, double& sah)
  {
    const BinNode& n = nodes[idx];
    if (!n.leaf()) {
      solve(n.left, sah);
      solve(n.right, sah);
    }
    solve_node(idx, sah);
  }

  void solve_node(int idx, double& sah)
  {
    const BinNode& n = nodes[idx];
    Entry& en = e[idx];
    std::memset(&en, 0, sizeof(en));
    const float area = n.box.half_area();
    sah += (n.leaf() ? (double)n.count : 1.0) * area;
    const float c_leaf = n.count <= kMaxLeaf ? area * (float)n.count * kPrimCost : FLT_MAX;
    if (n.leaf()) {
      for (int i = 1; i < 8; ++i)
        en.c[i] = c_leaf;
      en.leaf1 = true;
      return;
    }
    const Entry &l = e[n.left], &r = e[n.right];
    float dist[9];
    for (int j = 2; j <= 8; ++j) {
      dist[j] = FLT_MAX;
      for (int k = 1; k < j; ++k) {
        if (k > 7 || j - k > 7)
          continue;
        float c = l.c[k] + r.c[j - k];
        if (c < dist[j]) {
          dist[j] = c;
          en.split[j] = (uint8_t)k;
        }
      }
    }
    const float c_internal = dist[8] + area * kNodeCost;
    en.leaf1 = c_leaf <= c_internal;
    en.c[1] = en.leaf1 ? c_leaf : c_internal;
    for (int i = 2; i < 8; ++i) {
      en.use_dist[i] = dist[i] < en.c[i - 1];
      en.c[i] = en.use_dist[i] ? dist[i] : en.c[i - 1];
    }
  }

  bool is_leaf(int n) const { return e[n].leaf1; }

  // roots of the optimal forest of at most i trees for the subtree of n
  void gather(int n, int i, std::vector<int>& out) const
  {
    while (i > 1 && !(nodes[n].left >= 0 && e[n].use_dist[i]))
      --i;
    if (i == 1) {
      out.push_back(n);
      return;
    }
    const int k = e[n].split[i];
    gather(nodes[n].left, k, out);
    gather(nodes[n].right, i - k, out);
  }

  // children of the wide node that represents the (internal) subtree of n
  void gather_children(int n, std::vector<int>& out) const
  {
    const int k = e[n].split[8];
    gather(nodes[n].left, k, out);
    gather(nodes[n].right, 8 - k, out);
  }
};

// floor(log2) style exponent so that ext / 2^e <= 255, as a biased float exponent byte.
uint8_t
quant_exponent(float ext)
{
  if (!(ext > 0.f))
    return 1; // any tiny scale; q values will be 0
  int e;
  std::frexp(ext / 255.0f, &e); // ext/255 = m * 2^e, m in [0.5,1)  ->  2^e >= ext/255
  int biased = e + 127;
  if (biased < 1)
    biased = 1;
  if (biased > 254)
    biased = 254;
  return (uint8_t)biased;
}

// IEEE half-float bits of the integer q in 0..255 (exact).
uint16_t
half_of_byte(int q)
{
  if (q <= 0)
    return 0;
  int e = 0;
  while ((q >> (e + 1)) != 0)
    ++e; // q in [2^e, 2^(e+1))
  const int mant = (q - (1 << e)) << (10 - e);
  return (uint16_t)(((e + 15) << 10) | mant);
}

float
exp_scale(uint8_t e)
{
  uint32_t bits = (uint32_t)e << 23;
  float f;
  std::memcpy(&f, &bits, 4);
  return f;
}

// Treelet order (serial, deterministic post-pass over the breadth-first emission).  Breadth-first
// order keeps a node's children contiguous (required: child = child_base + rank) but puts every
// level of a subtree in a different region of the array, and the triangle records follow the node
// order; a ray walking down one subtree of a 1.4 GB structure touches one 2 MB page per level for
// the nodes and another for the triangles.  Here the nodes are re-emitted as treelets: starting at
// a subtree root, sibling groups are taken breadth-first until the treelet holds `budget` nodes;
// the groups left on the frontier start their own treelets right behind it, depth-first.  Sibling
// groups stay contiguous, every node's triangle block moves with it.  PTC_TREELET = nodes per
// treelet (0 keeps the breadth-first order).
static void
reorder_treelets(HostBvh& out)
{
  static const uint32_t budget = getenv("PTC_TREELET") ? (uint32_t)atoi(getenv("PTC_TREELET")) : 0u;
  const uint32_t n = (uint32_t)out.nodes.size();
  if (!budget || n < 2 * budget)
    return;
  std::vector<uint32_t> order; // new position -> old index
  order.reserve(n);
  order.push_back(0);
  std::vector<uint32_t> pending{ 0 }; // old indices whose child group starts a treelet (LIFO)
  std::vector<uint32_t> frontier, nextf;
  while (!pending.empty()) {
    const uint32_t root = pending.back();
    pending.pop_back();
    frontier.assign(1, root);
    uint32_t taken = 0;
    while (!frontier.empty()) {
      nextf.clear();
      size_t q = 0;
      for (; q < frontier.size(); ++q) {
        const Node8& nd = out.nodes[frontier[q]];
        const uint32_t k = (uint32_t)__builtin_popcount(nd.imask);
        if (!k)
          continue;
        if (taken && taken + k > budget)
          break;
        for (uint32_t c = 0; c < k; ++c) {
          order.push_back(nd.child_base + c);
          nextf.push_back(nd.child_base + c);
        }
        taken += k;
      }
      // what did not fit: roots of later treelets (reverse, so the first is handled next)
      for (size_t r = frontier.size(); r > q; --r)
        pending.push_back(frontier[r - 1]);
      if (q < frontier.size()) {
        for (size_t r = nextf.size(); r > 0; --r)
          pending.push_back(nextf[r - 1]);
        break;
      }
      frontier.swap(nextf);
    }
  }
  std::vector<uint32_t> new_of(n);
  for (uint32_t i = 0; i < n; ++i)
    new_of[order[i]] = i;
  decltype(out.nodes) nodes(n);
  decltype(out.tris) tris(out.tris.size());
  decltype(out.face_ids) ids(out.face_ids.size());
  uint32_t tri_top = 0;
  for (uint32_t i = 0; i < n; ++i) {
    Node8 nd = out.nodes[order[i]];
    if (nd.imask)
      nd.child_base = new_of[nd.child_base];
    const uint32_t nt = (uint32_t)__builtin_popcount(nd.valid24);
    for (uint32_t t = 0; t < nt; ++t) {
      tris[tri_top + t] = out.tris[nd.tri_base + t];
      ids[tri_top + t] = out.face_ids[nd.tri_base + t];
    }
    nd.tri_base = tri_top;
    tri_top += nt;
    nodes[i] = nd;
  }
  out.nodes.swap(nodes);
  out.tris.swap(tris);
  out.face_ids.swap(ids);
}

} // namespace

void
build_bvh(const ptc_scene_desc* desc, HostBvh& out)
{
  out = HostBvh();
  uint32_t n_faces = 0;
  for (uint32_t m = 0; m < desc->n_meshes; ++m)
    n_faces += desc->mesh_face_counts[m];

  if (n_faces == 0) {
    Node8 n;
    std::memset(&n, 0, sizeof(n));
    n.e[0] = n.e[1] = n.e[2] = 1;
    out.nodes.push_back(n);
    out.stats.wide_depth = 1;
    return;
  }

  const bool timing = getenv("PTC_BUILD_TIMING") != nullptr;
  auto now = [] { return std::chrono::steady_clock::now(); };
  auto secs = [](std::chrono::steady_clock::time_point a, std::chrono::steady_clock::time_point b) {
    return std::chrono::duration<double>(b - a).count();
  };
  const auto t_begin = now();
  // (a scene of a few thousand triangles builds in well under a millisecond on one thread; waking
  // workers would only add to it)
  Pool pool(n_faces >= 16384 ? build_threads() : 1u);
  const unsigned T = pool.size();
  Builder b(pool, n_faces);

  // Triangle boxes, padded.  The pad absorbs (a) the rounding of the slab test
  // against the exact box and (b) the fp32 error of Moeller-Trumbore near edges,
  // so a ray the reference's test accepts is never culled by a box.
  // (two threaded passes, the first over 1.12 GB of Face records at 10 M triangles; every
  // reduction is a min / max, so nothing depends on how the range is cut)
  Builder::Prim* const prims = b.buf[0].p;
  std::vector<Box> part(T), cpart(T);
  pool.chunks(0, n_faces, [&](uint32_t lo, uint32_t hi, unsigned k) {
    Box sb;
    for (uint32_t f = lo; f < hi; ++f) {
      Box fb;
      for (int v = 0; v < 3; ++v)
        fb.grow(desc->faces[f].vertices[v]);
      prims[f].box = fb;
      sb.grow(fb);
    }
    part[k] = sb;
  });
  Box scene;
  for (const Box& pb : part)
    if (pb.valid())
      scene.grow(pb);
  float extent = 0.f, maxabs = 0.f;
  for (int a = 0; a < 3; ++a) {
    extent = std::max(extent, scene.hi[a] - scene.lo[a]);
    maxabs = std::max(maxabs, std::max(std::fabs(scene.lo[a]), std::fabs(scene.hi[a])));
  }
  const float pad = std::ldexp(std::max(extent, maxabs), -15) + FLT_MIN;
  for (unsigned k = 0; k < T; ++k)
    part[k] = cpart[k] = Box();
  pool.chunks(0, n_faces, [&](uint32_t lo, uint32_t hi, unsigned k) {
    Box sb, cb;
    for (uint32_t f = lo; f < hi; ++f) {
      Builder::Prim& p = prims[f];
      for (int a = 0; a < 3; ++a) {
        p.box.lo[a] -= pad;
        p.box.hi[a] += pad;
        p.c[a] = 0.5f * (p.box.lo[a] + p.box.hi[a]);
      }
      p.id = f;
      sb.grow(p.box);
      cb.grow(p.c);
    }
    part[k] = sb, cpart[k] = cb;
  });
  Box root_box, root_cbox;
  for (unsigned k = 0; k < T; ++k)
    if (part[k].valid())
      root_box.grow(part[k]), root_cbox.grow(cpart[k]);

  auto emit_tri = [&](uint32_t at, uint32_t f) {
    const ptc_face& face = desc->faces[f];
    TriRecord t;
    std::memset(&t, 0, sizeof(t));
    for (int a = 0; a < 3; ++a) {
      t.v0[a] = face.vertices[0][a];
      t.e1[a] = face.vertices[1][a] - face.vertices[0][a]; // intersection.cuh:106
      t.e2[a] = face.vertices[2][a] - face.vertices[0][a]; // intersection.cuh:107
    }
    t.face_id = f;
    out.tris[at] = t;
    out.face_ids[at] = f;
  };

  const auto t0 = now();
  b.build(root_box, root_cbox);
  const int root = b.root;
  out.stats.max_depth = b.max_depth;
  const auto t1 = now();

  const CollapseDp dp(b, pool);
  {
    const float root_area = b.nodes[root].box.half_area();
    out.stats.sah_cost = root_area > 0 ? (float)(dp.sah_sum / root_area) : 0.f; // diagnostic
  }
  const auto t2 = now();

  // Emission, level by level (== the order a serial breadth-first queue would give, so node
  // and triangle indices do not depend on the thread count): per level, (1) every wide node
  // picks its children and their slots in parallel, (2) one scan hands out child_base /
  // tri_base, (3) the nodes and their triangles are written in parallel.
  struct Item
  {
    int bin;      // binary node whose subtree this wide node covers
    uint32_t out; // index in out.nodes
  };
  struct Plan
  {
    int kids[8];
    int kid_in_slot[8];
    int n_kids;
    uint32_t n_internal, n_tris, n_leaf_kids;
    uint32_t child_base, tri_base;
  };
  RawVector<Item> level, next;
  level.push_back(Item{ root, 0u });
  RawVector<Plan> plan;
  // (sized once, not value-initialised: the threads that write the records touch the pages)
  out.tris.resize(n_faces);
  out.face_ids.resize(n_faces);
  out.nodes.reserve(n_faces / 4 + 16);
  out.nodes.resize(1);
  uint32_t depth = 0, tri_top = 0;
  while (!level.empty()) {
    ++depth;
    plan.resize(level.size());
    // 1. children (the forest of binary subtrees the collapse DP chose) and slot assignment:
    //    slot s is visited first by rays whose direction is negative exactly on the axes set
    //    in s, so it should hold the child that comes first along sign_s.  Greedy on
    //    cost = centroid . sign_s.
    pool.chunks(0, (uint32_t)level.size(), [&](uint32_t lo, uint32_t hi, unsigned) {
      std::vector<int> kids;
      for (uint32_t q = lo; q < hi; ++q) {
        const int bin = level[q].bin;
        Plan& P = plan[q];
        kids.clear();
        if (dp.is_leaf(bin))
          kids.push_back(bin);
        else
          dp.gather_children(bin, kids);
        P.n_kids = (int)kids.size();
        const Box& nb = b.nodes[bin].box;
        float cen[3];
        for (int a = 0; a < 3; ++a)
          cen[a] = 0.5f * (nb.lo[a] + nb.hi[a]);
        int slot_of[8];
        std::fill(slot_of, slot_of + 8, -1);
        std::fill(P.kid_in_slot, P.kid_in_slot + 8, -1);
        float cost[8][8];
        for (int k = 0; k < P.n_kids; ++k) {
          P.kids[k] = kids[k];
          const Box& cb = b.nodes[kids[k]].box;
          for (int sl = 0; sl < 8; ++sl) {
            float c = 0.f;
            for (int a = 0; a < 3; ++a) {
              float d = 0.5f * (cb.lo[a] + cb.hi[a]) - cen[a];
              c += ((sl >> a) & 1) ? -d : d;
            }
            cost[k][sl] = c;
          }
        }
        for (int k = P.n_kids; k < 8; ++k)
          P.kids[k] = -1;
        for (int round = 0; round < P.n_kids; ++round) {
          float bc = FLT_MAX;
          int bk = -1, bs = -1;
          for (int k = 0; k < P.n_kids; ++k) {
            if (slot_of[k] >= 0)
              continue;
            for (int sl = 0; sl < 8; ++sl)
              if (P.kid_in_slot[sl] < 0 && cost[k][sl] < bc) {
                bc = cost[k][sl];
                bk = k;
                bs = sl;
              }
          }
          slot_of[bk] = bs;
          P.kid_in_slot[bs] = bk;
        }
        P.n_internal = P.n_tris = P.n_leaf_kids = 0;
        for (int sl = 0; sl < 8; ++sl) {
          const int k = P.kid_in_slot[sl];
          if (k < 0)
            continue;
          if (dp.is_leaf(P.kids[k])) {
            P.n_tris += b.nodes[P.kids[k]].count;
            P.n_leaf_kids++;
          } else
            P.n_internal++;
        }
      }
    });
    // 2. index hand-out in queue order
    uint32_t node_top = (uint32_t)out.nodes.size();
    for (Plan& P : plan) {
      P.child_base = node_top;
      P.tri_base = tri_top;
      node_top += P.n_internal;
      tri_top += P.n_tris;
      out.stats.n_leaves += P.n_leaf_kids;
    }
    const uint32_t next_first = (uint32_t)out.nodes.size();
    out.nodes.resize(node_top);
    next.resize(node_top - next_first);
    // 3. emit
    pool.chunks(0, (uint32_t)level.size(), [&](uint32_t lo, uint32_t hi, unsigned) {
      for (uint32_t q = lo; q < hi; ++q) {
        const Plan& P = plan[q];
        const Box& nb = b.nodes[level[q].bin].box;
        Node8 n;
        std::memset(&n, 0, sizeof(n));
        for (int a = 0; a < 3; ++a) {
          n.origin[a] = nb.lo[a];
          n.e[a] = quant_exponent(nb.hi[a] - nb.lo[a]);
        }
        n.child_base = P.child_base;
        n.tri_base = P.tri_base;
        uint32_t child = P.child_base, tri = P.tri_base;
        for (int sl = 0; sl < 8; ++sl) {
          const int k = P.kid_in_slot[sl];
          if (k < 0) { // empty slot: an inverted box (lo > hi) can never be hit
            for (int a = 0; a < 3; ++a) {
              n.lo_q(a, sl) = half_of_byte(255);
              n.hi_q(a, sl) = half_of_byte(0);
            }
            continue;
          }
          n.occupied |= 1u << sl;
          const BinNode& c = b.nodes[P.kids[k]];
          for (int a = 0; a < 3; ++a) {
            float sc = exp_scale(n.e[a]);
            float ql = std::floor((c.box.lo[a] - n.origin[a]) / sc);
            float qh = std::ceil((c.box.hi[a] - n.origin[a]) / sc);
            // guard against the division rounding the wrong way
            while (ql > 0.f && n.origin[a] + ql * sc > c.box.lo[a])
              ql -= 1.f;
            while (qh < 255.f && n.origin[a] + qh * sc < c.box.hi[a])
              qh += 1.f;
            n.lo_q(a, sl) = half_of_byte((int)std::min(255.f, std::max(0.f, ql)));
            n.hi_q(a, sl) = half_of_byte((int)std::min(255.f, std::max(0.f, qh)));
          }
          if (dp.is_leaf(P.kids[k])) {
            n.valid24 |= ((1u << c.count) - 1u) << (3 * sl); // 001, 011, 111
            uint32_t ids[kMaxLeaf];
            for (uint32_t t = 0; t < c.count; ++t)
              ids[t] = b.ids[c.first + t];
            for (uint32_t t = 1; t < c.count; ++t) // ascending face id (<= 3 entries)
              for (uint32_t u = t; u > 0 && ids[u] < ids[u - 1]; --u)
                std::swap(ids[u], ids[u - 1]);
            for (uint32_t t = 0; t < c.count; ++t)
              emit_tri(tri++, ids[t]);
          } else {
            n.imask |= (uint8_t)(1u << sl);
            next[child - next_first] = Item{ P.kids[k], child };
            ++child;
          }
        }
        out.nodes[level[q].out] = n;
      }
    });
    level.swap(next);
  }
  out.tris.resize(tri_top); // == n_faces: every face is emitted exactly once
  out.face_ids.resize(tri_top);
  out.stats.wide_depth = depth;
  reorder_treelets(out);
  if (timing)
    fprintf(stderr, "build_bvh: %u faces, %u threads: boxes %.2f s, binary SAH %.2f s (%u wide nodes, %zu tasks), collapse DP %.2f s, emit %.2f s\n",
            n_faces, T, secs(t_begin, t0), secs(t0, t1), b.n_wide, b.tasks.size(), secs(t1, t2), secs(t2, now()));
}

uint32_t
build_shade_records(const ptc_scene_desc* desc, RawVector<ShadeRecord>& out)
{
  std::atomic<uint32_t> max_material{ 0 };
  uint32_t n_faces = 0;
  for (uint32_t m = 0; m < desc->n_meshes; ++m)
    n_faces += desc->mesh_face_counts[m];
  out.resize(n_faces);
  parallel_for(n_faces, [&](uint32_t lo, uint32_t hi) {
    uint32_t mx = 0;
    for (uint32_t f = lo; f < hi; ++f) {
      const ptc_face& face = desc->faces[f];
      ShadeRecord& s = out[f];
      std::memcpy(s.n, face.normals, sizeof(s.n));
      std::memcpy(s.uv, face.texcoords, sizeof(s.uv));
      std::memcpy(s.tangent, face.tangent, sizeof(s.tangent));
      s.material_id = face.material_id;
      mx = std::max(mx, (uint32_t)face.material_id);
    }
    uint32_t seen = max_material.load();
    while (mx > seen && !max_material.compare_exchange_weak(seen, mx)) {
    }
  });
  return max_material.load();
}

} // namespace ptc

// ---- host-only C ABI (include/ptcore.h) -------------------------------------
extern "C" int
ptc_host_bvh_build(const ptc_scene_desc* desc, ptc_host_bvh* out)
{
  if (!desc || !out)
    return PTC_ERR_INVALID;
  ptc::HostBvh* h = new ptc::HostBvh();
  ptc::build_bvh(desc, *h);
  out->nodes = h->nodes.data();
  out->n_nodes = (uint32_t)h->nodes.size();
  out->node_bytes = sizeof(ptc::Node8);
  out->tris = h->tris.data();
  out->n_tris = (uint32_t)h->tris.size();
  out->tri_bytes = sizeof(ptc::TriRecord);
  out->tri_face_ids = h->face_ids.data();
  out->sah_cost = h->stats.sah_cost;
  out->max_depth = h->stats.max_depth;
  out->_owner = h;
  return PTC_OK;
}

extern "C" void
ptc_host_bvh_free(ptc_host_bvh* bvh)
{
  if (bvh && bvh->_owner) {
    delete (ptc::HostBvh*)bvh->_owner;
    bvh->_owner = nullptr;
  }
}
