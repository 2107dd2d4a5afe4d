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
#include <cfloat>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <numeric>

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
  uint32_t first = 0, count = 0; // ... or a range of `order` (leaf)
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


struct Builder
{
  const std::vector<Box>& boxes;
  std::vector<float> centroid; // 3 per primitive
  std::vector<uint32_t> order;
  std::vector<BinNode> nodes;
  uint32_t max_depth = 0;

  explicit Builder(const std::vector<Box>& b) : boxes(b)
  {
    centroid.resize(b.size() * 3);
    order.resize(b.size());
    std::iota(order.begin(), order.end(), 0u);
    for (size_t i = 0; i < b.size(); ++i)
      for (int a = 0; a < 3; ++a)
        centroid[3 * i + a] = 0.5f * (b[i].lo[a] + b[i].hi[a]);
    nodes.reserve(b.size() * 2 + 1);
  }

  int build(uint32_t first, uint32_t count, uint32_t depth)
  {
    max_depth = std::max(max_depth, depth);
    int idx = (int)nodes.size();
    nodes.emplace_back();
    Box box, cbox;
    for (uint32_t i = first; i < first + count; ++i) {
      box.grow(boxes[order[i]]);
      cbox.grow(&centroid[3 * order[i]]);
    }
    nodes[idx].box = box;
    nodes[idx].first = first;
    nodes[idx].count = count;
    if (count <= 1)
      return idx;

    // binned SAH over the three axes
    float best_cost = FLT_MAX;
    int best_axis = -1, best_split = -1;
    for (int a = 0; a < 3; ++a) {
      float ext = cbox.hi[a] - cbox.lo[a];
      if (!(ext > 0.f))
        continue;
      Box bin_box[kBins];
      uint32_t bin_cnt[kBins] = { 0 };
      float scale = kBins / ext;
      for (uint32_t i = first; i < first + count; ++i) {
        uint32_t p = order[i];
        int b = std::min(kBins - 1, (int)((centroid[3 * p + a] - cbox.lo[a]) * scale));
        bin_box[b].grow(boxes[p]);
        bin_cnt[b]++;
      }
      float right_area[kBins];
      uint32_t right_cnt[kBins];
      Box acc;
      uint32_t cnt = 0;
      for (int b = kBins - 1; b > 0; --b) {
        if (bin_cnt[b])
          acc.grow(bin_box[b]);
        cnt += bin_cnt[b];
        right_area[b] = acc.valid() ? acc.half_area() : 0.f;
        right_cnt[b] = cnt;
      }
      acc = Box();
      cnt = 0;
      for (int b = 0; b < kBins - 1; ++b) {
        if (bin_cnt[b])
          acc.grow(bin_box[b]);
        cnt += bin_cnt[b];
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
      float ext = cbox.hi[best_axis] - cbox.lo[best_axis];
      float scale = kBins / ext;
      float lo = cbox.lo[best_axis];
      auto it = std::partition(order.begin() + first, order.begin() + first + count, [&](uint32_t p) {
        int b = std::min(kBins - 1, (int)((centroid[3 * p + best_axis] - lo) * scale));
        return b <= best_split;
      });
      mid = (uint32_t)(it - order.begin());
      if (mid == first || mid == first + count)
        mid = first + count / 2;
    }
    int l = build(first, mid - first, depth + 1);
    int r = build(mid, first + count - mid, depth + 1);
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
  std::vector<Entry> e;

  explicit CollapseDp(const std::vector<BinNode>& n) : nodes(n), e(n.size())
  {
    for (int idx = (int)nodes.size() - 1; idx >= 0; --idx) { // children have larger indices
      const BinNode& n = nodes[idx];
      Entry& en = e[idx];
      std::memset(&en, 0, sizeof(en));
      const float area = n.box.half_area();
      const float c_leaf = n.count <= kMaxLeaf ? area * (float)n.count * kPrimCost : FLT_MAX;
      if (n.leaf()) {
        for (int i = 1; i < 8; ++i)
          en.c[i] = c_leaf;
        en.leaf1 = true;
        continue;
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
  std::vector<Box> boxes(n_faces);
  Box scene;
  for (uint32_t f = 0; f < n_faces; ++f) {
    for (int v = 0; v < 3; ++v)
      boxes[f].grow(desc->faces[f].vertices[v]);
    scene.grow(boxes[f]);
  }
  float extent = 0.f, maxabs = 0.f;
  if (n_faces)
    for (int a = 0; a < 3; ++a) {
      extent = std::max(extent, scene.hi[a] - scene.lo[a]);
      maxabs = std::max(maxabs, std::max(std::fabs(scene.lo[a]), std::fabs(scene.hi[a])));
    }
  const float pad = std::ldexp(std::max(extent, maxabs), -15) + FLT_MIN;
  for (Box& b : boxes)
    for (int a = 0; a < 3; ++a) {
      b.lo[a] -= pad;
      b.hi[a] += pad;
    }

  auto emit_tri = [&](uint32_t f) {
    const ptc_face& face = desc->faces[f];
    TriRecord t;
    std::memset(&t, 0, sizeof(t));
    for (int a = 0; a < 3; ++a) {
      t.v0[a] = face.vertices[0][a];
      t.e1[a] = face.vertices[1][a] - face.vertices[0][a]; // intersection.cuh:106
      t.e2[a] = face.vertices[2][a] - face.vertices[0][a]; // intersection.cuh:107
    }
    t.face_id = f;
    out.tris.push_back(t);
    out.face_ids.push_back(f);
  };

  if (n_faces == 0) {
    Node8 n;
    std::memset(&n, 0, sizeof(n));
    n.e[0] = n.e[1] = n.e[2] = 1;
    out.nodes.push_back(n);
    out.stats.wide_depth = 1;
    return;
  }

  Builder b(boxes);
  int root = b.build(0, n_faces, 0);
  out.stats.max_depth = b.max_depth;

  // SAH cost of the binary tree (diagnostic)
  {
    float root_area = b.nodes[root].box.half_area();
    double cost = 0;
    for (const BinNode& n : b.nodes)
      cost += (n.leaf() ? (double)n.count : 1.0) * n.box.half_area();
    out.stats.sah_cost = root_area > 0 ? (float)(cost / root_area) : 0.f;
  }

  const CollapseDp dp(b.nodes);

  struct Pending
  {
    int bin;      // binary node whose subtree this wide node covers
    uint32_t out; // index in out.nodes
    uint32_t depth;
  };
  std::vector<Pending> queue;
  out.nodes.emplace_back();
  queue.push_back({ root, 0, 1 });

  for (size_t qi = 0; qi < queue.size(); ++qi) {
    const Pending cur = queue[qi];
    const BinNode& top = b.nodes[cur.bin];
    out.stats.wide_depth = std::max(out.stats.wide_depth, cur.depth);

    // 1. gather up to 8 children: the forest of binary subtrees the collapse DP chose
    std::vector<int> kids;
    if (dp.is_leaf(cur.bin))
      kids.push_back(cur.bin);
    else
      dp.gather_children(cur.bin, kids);

    // 2. slot assignment: slot s is visited first by rays whose direction is
    //    negative exactly on the axes set in s, so it should hold the child that
    //    comes first along sign_s.  Greedy on cost = centroid . sign_s.
    Box nb = top.box;
    float cen[3];
    for (int a = 0; a < 3; ++a)
      cen[a] = 0.5f * (nb.lo[a] + nb.hi[a]);
    int slot_of[8];
    int kid_in_slot[8];
    std::fill(slot_of, slot_of + 8, -1);
    std::fill(kid_in_slot, kid_in_slot + 8, -1);
    float cost[8][8];
    for (size_t k = 0; k < kids.size(); ++k) {
      const Box& cb = b.nodes[kids[k]].box;
      for (int s = 0; s < 8; ++s) {
        float c = 0.f;
        for (int a = 0; a < 3; ++a) {
          float d = 0.5f * (cb.lo[a] + cb.hi[a]) - cen[a];
          c += ((s >> a) & 1) ? -d : d;
        }
        cost[k][s] = c;
      }
    }
    for (size_t round = 0; round < kids.size(); ++round) {
      float bc = FLT_MAX;
      int bk = -1, bs = -1;
      for (size_t k = 0; k < kids.size(); ++k) {
        if (slot_of[k] >= 0)
          continue;
        for (int s = 0; s < 8; ++s)
          if (kid_in_slot[s] < 0 && cost[k][s] < bc) {
            bc = cost[k][s];
            bk = (int)k;
            bs = s;
          }
      }
      slot_of[bk] = bs;
      kid_in_slot[bs] = bk;
    }

    // 3. emit
    Node8 n;
    std::memset(&n, 0, sizeof(n));
    for (int a = 0; a < 3; ++a) {
      n.origin[a] = nb.lo[a];
      n.e[a] = quant_exponent(nb.hi[a] - nb.lo[a]);
    }
    n.child_base = (uint32_t)out.nodes.size();
    n.tri_base = (uint32_t)out.tris.size();
    for (int s = 0; s < 8; ++s) {
      int k = kid_in_slot[s];
      if (k < 0) { // empty slot: an inverted box (lo > hi) can never be hit
        for (int a = 0; a < 3; ++a) {
          n.lo[a][s] = half_of_byte(255);
          n.hi[a][s] = half_of_byte(0);
        }
        continue;
      }
      n.occupied |= 1u << s;
      const BinNode& c = b.nodes[kids[k]];
      for (int a = 0; a < 3; ++a) {
        float sc = exp_scale(n.e[a]);
        float ql = std::floor((c.box.lo[a] - n.origin[a]) / sc);
        float qh = std::ceil((c.box.hi[a] - n.origin[a]) / sc);
        // guard against the division rounding the wrong way
        while (ql > 0.f && n.origin[a] + ql * sc > c.box.lo[a])
          ql -= 1.f;
        while (qh < 255.f && n.origin[a] + qh * sc < c.box.hi[a])
          qh += 1.f;
        n.lo[a][s] = half_of_byte((int)std::min(255.f, std::max(0.f, ql)));
        n.hi[a][s] = half_of_byte((int)std::min(255.f, std::max(0.f, qh)));
      }
      if (dp.is_leaf(kids[k])) {
        n.valid24 |= ((1u << c.count) - 1u) << (3 * s); // 001, 011, 111
        std::vector<uint32_t> ids(b.order.begin() + c.first, b.order.begin() + c.first + c.count);
        std::sort(ids.begin(), ids.end());
        for (uint32_t f : ids)
          emit_tri(f);
        out.stats.n_leaves++;
      } else {
        n.imask |= (uint8_t)(1u << s);
        queue.push_back({ kids[k], (uint32_t)out.nodes.size(), cur.depth + 1 });
        out.nodes.emplace_back();
      }
    }
    out.nodes[cur.out] = n;
  }
}

void
build_shade_records(const ptc_scene_desc* desc, std::vector<ShadeRecord>& out)
{
  uint32_t n_faces = 0;
  for (uint32_t m = 0; m < desc->n_meshes; ++m)
    n_faces += desc->mesh_face_counts[m];
  out.resize(n_faces);
  for (uint32_t f = 0; f < n_faces; ++f) {
    const ptc_face& face = desc->faces[f];
    ShadeRecord& s = out[f];
    std::memcpy(s.n, face.normals, sizeof(s.n));
    std::memcpy(s.uv, face.texcoords, sizeof(s.uv));
    std::memcpy(s.tangent, face.tangent, sizeof(s.tangent));
    s.material_id = face.material_id;
  }
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
