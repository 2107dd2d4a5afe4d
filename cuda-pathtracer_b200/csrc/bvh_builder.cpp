// bvh_builder.cpp -- host builder for the 8-wide compressed BVH (bvh.h).
//
// Pipeline: padded triangle boxes -> binary BVH by binned SAH (leaves <= 3
// triangles) -> collapse to 8-wide by repeatedly opening the child with the
// largest surface area -> octant-ordered slot assignment -> breadth-first
// emission (internal children contiguous, a node's triangles contiguous) with
// outward-rounded 8-bit quantisation.
//
// New work relative to the reference, which tests every face per ray
// (cuda_opengl/include/shaders/intersection.cuh:179-196).
#include "bvh.h"

#include <algorithm>
#include <atomic>
#include <cfloat>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <numeric>
#include <thread>

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
  uint32_t first = 0, count = 0; // ... or a range of the partitioned primitive array (leaf)
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

// Binned-SAH binary tree, built top-down.  Parallel in two ways (scene.cpp:202-283 builds
// nothing, so at 10 M triangles this step IS the load time): the bin/partition passes of a
// large node are split over the worker threads, and the two children of a node with more than
// kSpawn primitives are built concurrently.  Node slots come from one atomic counter, so a
// child always has a larger index than its parent (the collapse DP walks indices downwards);
// the tree itself does not depend on the thread count or on timing -- every reduction is a
// min/max or an integer sum.
struct Builder
{
  static constexpr uint32_t kSpawn = 1u << 16;    // children of bigger nodes run as tasks
  static constexpr uint32_t kParallel = 1u << 19; // bigger nodes bin their range in parallel

  // Primitive records are PARTITIONED IN PLACE (no index indirection): every pass over a node's
  // range is a sequential read, which is what makes the passes scale across threads.
  struct Prim
  {
    Box box;
    float c[3]; // centroid
    uint32_t id;
  };
  std::vector<Prim> prims;
  // 2 n + 1 node slots, handed out by n_nodes.  NOT value-initialised up front: constructing 20 M
  // BinNodes (0.8 GB) on one thread was a third of the SAH phase at 10 M triangles; a slot is
  // initialised by the build() call that takes it, on the thread that then fills it.
  struct RawNodes
  {
    BinNode* p = nullptr;
    size_t n = 0;
    ~RawNodes() { std::free(p); }
    void reserve(size_t count)
    {
      p = static_cast<BinNode*>(std::malloc(count * sizeof(BinNode)));
      n = count;
    }
    BinNode& operator[](size_t i) { return p[i]; }
    const BinNode& operator[](size_t i) const { return p[i]; }
  } raw;
  std::vector<BinNode> nodes; // the used prefix, copied out after the build (finish)
  std::atomic<uint32_t> n_nodes{ 0 };
  std::atomic<uint32_t> max_depth{ 0 };
  std::atomic<int> idle_threads;
  const unsigned n_threads;

  explicit Builder(const std::vector<Box>& b) : idle_threads(0), n_threads(build_threads())
  {
    idle_threads = (int)n_threads - 1;
    prims.resize(b.size());
    parallel_chunks(0, (uint32_t)b.size(), true, [&](uint32_t lo, uint32_t hi, unsigned) {
      for (uint32_t i = lo; i < hi; ++i) {
        prims[i].box = b[i];
        prims[i].id = i;
        for (int a = 0; a < 3; ++a)
          prims[i].c[a] = 0.5f * (b[i].lo[a] + b[i].hi[a]);
      }
    });
    raw.reserve(b.size() * 2 + 1);
  }

  // after build(): the used slots as a vector (what the collapse and the emission read)
  void finish()
  {
    const uint32_t n = n_nodes.load();
    nodes.resize(n);
    BinNode* dst = nodes.data();
    const BinNode* src = raw.p;
    parallel_for(n, [=](uint32_t lo, uint32_t hi) { std::memcpy(dst + lo, src + lo, (size_t)(hi - lo) * sizeof(BinNode)); });
    std::free(raw.p);
    raw.p = nullptr;
  }

  // f(lo, hi, chunk) over [first, last) in up to n_threads contiguous chunks
  template <typename F>
  void parallel_chunks(uint32_t first, uint32_t last, bool parallel, F f)
  {
    const uint32_t n = last - first;
    const unsigned t = parallel && n >= kParallel ? n_threads : 1u;
    if (t <= 1) {
      f(first, last, 0u);
      return;
    }
    std::vector<std::thread> th;
    const uint32_t per = (n + t - 1) / t;
    for (unsigned k = 1; k < t; ++k) {
      const uint32_t lo = first + std::min(n, k * per), hi = first + std::min(n, (k + 1) * per);
      if (lo < hi)
        th.emplace_back([=, &f] { f(lo, hi, k); });
    }
    f(first, first + std::min(n, per), 0u);
    for (std::thread& x : th)
      x.join();
  }

  struct Bins
  {
    Box box[3][kBins];
    uint32_t cnt[3][kBins];
    Bins() { std::memset(cnt, 0, sizeof(cnt)); }
  };

  int build(uint32_t first, uint32_t count, uint32_t depth, bool top = true)
  {
    uint32_t seen = max_depth.load(std::memory_order_relaxed);
    while (depth > seen && !max_depth.compare_exchange_weak(seen, depth, std::memory_order_relaxed)) {
    }
    const int idx = (int)n_nodes.fetch_add(1u);
    new (&raw[idx]) BinNode();
    BinNode* const nodes = raw.p; // this function fills raw slots; the member vector is built by finish()
    // `top`: no other task is running beside this node, its passes may use all the threads
    const bool wide = top && count >= kParallel && n_threads > 1;

    // (the per-chunk partial results of a node that runs on one thread live on the stack: two
    // heap allocations per node, 20 M nodes, were a fifth of the single-thread build)
    Box box, cbox;
    if (!wide) {
      for (uint32_t i = first; i < first + count; ++i) {
        box.grow(prims[i].box);
        cbox.grow(prims[i].c);
      }
    } else {
      std::vector<Box> pb(n_threads), pc(n_threads);
      parallel_chunks(first, first + count, wide, [&](uint32_t lo, uint32_t hi, unsigned k) {
        Box bb, cc;
        for (uint32_t i = lo; i < hi; ++i) {
          bb.grow(prims[i].box);
          cc.grow(prims[i].c);
        }
        pb[k] = bb, pc[k] = cc;
      });
      for (size_t k = 0; k < pb.size(); ++k)
        if (pb[k].valid())
          box.grow(pb[k]), cbox.grow(pc[k]);
    }
    nodes[idx].box = box;
    nodes[idx].first = first;
    nodes[idx].count = count;
    if (count <= 1)
      return idx;

    // binned SAH over the three axes (one pass fills the bins of all three)
    float ext[3], scale[3];
    for (int a = 0; a < 3; ++a) {
      ext[a] = cbox.hi[a] - cbox.lo[a];
      scale[a] = ext[a] > 0.f ? kBins / ext[a] : 0.f;
    }
    Bins one;
    std::vector<Bins> many(wide ? n_threads : 0u);
    Bins* const part = wide ? many.data() : &one;
    const size_t n_part = wide ? many.size() : 1u;
    parallel_chunks(first, first + count, wide, [&](uint32_t lo, uint32_t hi, unsigned k) {
      Bins& B = part[k];
      for (uint32_t i = lo; i < hi; ++i) {
        const Prim& p = prims[i];
        for (int a = 0; a < 3; ++a) {
          if (!(ext[a] > 0.f))
            continue;
          const int bn = std::min(kBins - 1, (int)((p.c[a] - cbox.lo[a]) * scale[a]));
          B.box[a][bn].grow(p.box);
          B.cnt[a][bn]++;
        }
      }
    });
    Bins& B = part[0];
    for (size_t k = 1; k < n_part; ++k)
      for (int a = 0; a < 3; ++a)
        for (int bn = 0; bn < kBins; ++bn)
          if (part[k].cnt[a][bn]) {
            B.box[a][bn].grow(part[k].box[a][bn]);
            B.cnt[a][bn] += part[k].cnt[a][bn];
          }

    float best_cost = FLT_MAX;
    int best_axis = -1, best_split = -1;
    for (int a = 0; a < 3; ++a) {
      if (!(ext[a] > 0.f))
        continue;
      float right_area[kBins];
      uint32_t right_cnt[kBins];
      Box acc;
      uint32_t cnt = 0;
      for (int b = kBins - 1; b > 0; --b) {
        if (B.cnt[a][b])
          acc.grow(B.box[a][b]);
        cnt += B.cnt[a][b];
        right_area[b] = acc.valid() ? acc.half_area() : 0.f;
        right_cnt[b] = cnt;
      }
      acc = Box();
      cnt = 0;
      for (int b = 0; b < kBins - 1; ++b) {
        if (B.cnt[a][b])
          acc.grow(B.box[a][b]);
        cnt += B.cnt[a][b];
        if (!cnt || !right_cnt[b + 1])
          continue;
        float cost = acc.half_area() * cnt + right_area[b + 1] * right_cnt[b + 1];
        if (cost < best_cost) {
          best_cost = cost;
          best_axis = a;
          best_split = b;
        }
      }
    }

    uint32_t mid;
    if (best_axis < 0) {
      // all centroids coincide: leaf if small enough, else split in the middle
      if (count <= kMaxLeaf)
        return idx;
      mid = first + count / 2;
    } else {
      float leaf_cost = box.half_area() * count;
      if (count <= kMaxLeaf && leaf_cost <= best_cost + box.half_area() * trav_cost())
        return idx; // cheaper (or as cheap) to keep as a leaf
      const float sc = scale[best_axis];
      const float lo = cbox.lo[best_axis];
      auto it = std::partition(prims.begin() + first, prims.begin() + first + count, [&](const Prim& p) {
        int b = std::min(kBins - 1, (int)((p.c[best_axis] - lo) * sc));
        return b <= best_split;
      });
      mid = (uint32_t)(it - prims.begin());
      if (mid == first || mid == first + count)
        mid = first + count / 2;
    }

    int l, r;
    const uint32_t nl = mid - first, nr = first + count - mid;
    // both children big: build the left one on another thread if one is free.  While a node
    // is still "top" only one of its children can stay so (the larger), the other runs beside it.
    bool spawned = false;
    if (std::min(nl, nr) >= kSpawn && n_threads > 1) {
      int idle = idle_threads.load();
      while (idle > 0 && !idle_threads.compare_exchange_weak(idle, idle - 1)) {
      }
      spawned = idle > 0;
    }
    if (spawned) {
      std::thread t([&] { l = build(first, nl, depth + 1, false); });
      r = build(mid, nr, depth + 1, false);
      t.join();
      idle_threads.fetch_add(1);
    } else {
      const bool keep_top = top && std::min(nl, nr) < kSpawn;
      l = build(first, nl, depth + 1, keep_top && nl >= nr);
      r = build(mid, nr, depth + 1, keep_top && nr > nl);
    }
    nodes[idx].left = l;
    nodes[idx].right = r;
    return idx;
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
  const std::vector<BinNode>& nodes;
  struct Entry
  {
    float c[8];        // c[i], i = 1..7
    uint8_t split[9];  // distribute(j): roots given to the left child, j = 2..8
    uint8_t use_dist[8]; // c[i] came from distribute(i) (else from c[i-1]), i = 2..7
    bool leaf1;        // c[1] is the leaf alternative
  };
  std::unique_ptr<Entry[]> e; // one per binary node; every entry is written by solve()
  std::atomic<int> idle_threads;
  static constexpr uint32_t kSpawn = 1u << 15; // subtrees with more primitives may run as tasks

  // An entry depends on its two children only, so the subtrees of a node are independent: big
  // ones run concurrently (the serial index-order loop this replaces was 0.5 s at 10 M triangles).
  explicit CollapseDp(const std::vector<BinNode>& n, int root)
    : nodes(n), e(new Entry[n.size()]), idle_threads((int)build_threads() - 1)
  {
    solve(root);
  }

  void solve(int idx)
  {
    {
      const BinNode& n = nodes[idx];
      if (!n.leaf()) {
        bool spawned = false;
        if (n.count >= kSpawn && nodes[n.left].count >= kSpawn / 2 && nodes[n.right].count >= kSpawn / 2) {
          int idle = idle_threads.load();
          while (idle > 0 && !idle_threads.compare_exchange_weak(idle, idle - 1)) {
          }
          spawned = idle > 0;
        }
        if (spawned) {
          std::thread t([&] { solve(n.left); });
          solve(n.right);
          t.join();
          idle_threads.fetch_add(1);
        } else {
          solve(n.left);
          solve(n.right);
        }
      }
    }
    {
      const BinNode& n = nodes[idx];
      Entry& en = e[idx];
      std::memset(&en, 0, sizeof(en));
      const float area = n.box.half_area();
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
  std::vector<Node8> nodes(n);
  std::vector<TriRecord> tris(out.tris.size());
  std::vector<uint32_t> ids(out.face_ids.size());
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

  // Triangle boxes, padded.  The pad absorbs (a) the rounding of the slab test
  // against the exact box and (b) the fp32 error of Moeller-Trumbore near edges,
  // so a ray the reference's test accepts is never culled by a box.
  const bool timing = getenv("PTC_BUILD_TIMING") != nullptr;
  auto now = [] { return std::chrono::steady_clock::now(); };
  auto secs = [](std::chrono::steady_clock::time_point a, std::chrono::steady_clock::time_point b) {
    return std::chrono::duration<double>(b - a).count();
  };
  const auto t_begin = now();
  // (threaded like everything below: one pass over 1.12 GB of Face records at 10 M triangles; the
  // scene box is a min/max reduction, so it does not depend on how the range is cut)
  std::vector<Box> boxes(n_faces);
  Box scene;
  {
    const unsigned t = build_threads();
    std::vector<Box> part(t);
    const uint32_t per = (n_faces + t - 1) / t;
    parallel_for(n_faces, [&](uint32_t lo, uint32_t hi) {
      Box sb;
      for (uint32_t f = lo; f < hi; ++f) {
        for (int v = 0; v < 3; ++v)
          boxes[f].grow(desc->faces[f].vertices[v]);
        sb.grow(boxes[f]);
      }
      part[per ? lo / per : 0] = sb; // parallel_for cuts [0, n) into chunks of `per`
    });
    for (const Box& pb : part)
      if (pb.valid())
        scene.grow(pb);
  }
  float extent = 0.f, maxabs = 0.f;
  if (n_faces)
    for (int a = 0; a < 3; ++a) {
      extent = std::max(extent, scene.hi[a] - scene.lo[a]);
      maxabs = std::max(maxabs, std::max(std::fabs(scene.lo[a]), std::fabs(scene.hi[a])));
    }
  const float pad = std::ldexp(std::max(extent, maxabs), -15) + FLT_MIN;
  parallel_for(n_faces, [&](uint32_t lo, uint32_t hi) {
    for (uint32_t f = lo; f < hi; ++f)
      for (int a = 0; a < 3; ++a) {
        boxes[f].lo[a] -= pad;
        boxes[f].hi[a] += pad;
      }
  });

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

  if (n_faces == 0) {
    Node8 n;
    std::memset(&n, 0, sizeof(n));
    n.e[0] = n.e[1] = n.e[2] = 1;
    out.nodes.push_back(n);
    out.stats.wide_depth = 1;
    return;
  }

  const auto t0 = now();
  if (timing)
    fprintf(stderr, "build_bvh: triangle boxes %.2f s\n", secs(t_begin, t0));
  Builder b(boxes);
  if (timing)
    fprintf(stderr, "build_bvh: builder set-up %.2f s\n", secs(t0, now()));
  int root = b.build(0, n_faces, 0);
  b.finish();
  out.stats.max_depth = b.max_depth.load();
  const auto t1 = now();

  // SAH cost of the binary tree (diagnostic)
  {
    float root_area = b.nodes[root].box.half_area();
    double cost = 0;
    for (const BinNode& n : b.nodes)
      cost += (n.leaf() ? (double)n.count : 1.0) * n.box.half_area();
    out.stats.sah_cost = root_area > 0 ? (float)(cost / root_area) : 0.f;
  }

  const CollapseDp dp(b.nodes, root);
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
    uint32_t n_internal, n_tris;
    uint32_t child_base, tri_base;
  };
  std::vector<Item> level{ Item{ root, 0u } }, next;
  std::vector<Plan> plan;
  out.tris.reserve(n_faces);     // every face is emitted exactly once: no regrowth copies
  out.face_ids.reserve(n_faces);
  out.nodes.reserve(n_faces / 4 + 16);
  out.nodes.resize(1);
  uint32_t depth = 0;
  while (!level.empty()) {
    ++depth;
    plan.resize(level.size());
    // 1. children (the forest of binary subtrees the collapse DP chose) and slot assignment:
    //    slot s is visited first by rays whose direction is negative exactly on the axes set
    //    in s, so it should hold the child that comes first along sign_s.  Greedy on
    //    cost = centroid . sign_s.
    parallel_for((uint32_t)level.size(), [&](uint32_t lo, uint32_t hi) {
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
        P.n_internal = P.n_tris = 0;
        for (int sl = 0; sl < 8; ++sl) {
          const int k = P.kid_in_slot[sl];
          if (k < 0)
            continue;
          if (dp.is_leaf(P.kids[k]))
            P.n_tris += b.nodes[P.kids[k]].count;
          else
            P.n_internal++;
        }
      }
    });
    // 2. index hand-out in queue order
    uint32_t node_top = (uint32_t)out.nodes.size(), tri_top = (uint32_t)out.tris.size();
    for (Plan& P : plan) {
      P.child_base = node_top;
      P.tri_base = tri_top;
      node_top += P.n_internal;
      tri_top += P.n_tris;
    }
    const uint32_t next_first = (uint32_t)out.nodes.size();
    out.nodes.resize(node_top);
    out.tris.resize(tri_top);
    out.face_ids.resize(tri_top);
    next.resize(node_top - next_first);
    // 3. emit
    parallel_for((uint32_t)level.size(), [&](uint32_t lo, uint32_t hi) {
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
              ids[t] = b.prims[c.first + t].id;
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
    for (const Plan& P : plan)
      for (int sl = 0; sl < 8; ++sl)
        if (P.kid_in_slot[sl] >= 0 && dp.is_leaf(P.kids[P.kid_in_slot[sl]]))
          out.stats.n_leaves++;
    level.swap(next);
  }
  out.stats.wide_depth = depth;
  reorder_treelets(out);
  if (timing)
    fprintf(stderr, "build_bvh: %u faces: binary SAH %.2f s, collapse DP %.2f s, emit %.2f s\n", n_faces,
            secs(t0, t1), secs(t1, t2), secs(t2, now()));
}

uint32_t
build_shade_records(const ptc_scene_desc* desc, std::vector<ShadeRecord>& out)
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
