// bz_bvh.cpp -- host-side build of the exact cull hierarchy used by BZ_MODE_CULLED.
//
// The reference tests every patch for every ray (bezierMesh.cpp:207 "Brute-force for now"); a patch
// reaches the root search only if the ray meets its FLAT underlying triangle (control points 0, 1, 2)
// in front of the start (bezierTriangle.cpp:124-131).  A 4-wide bounding-volume hierarchy over the
// axis-aligned boxes of those flat triangles therefore preserves the candidate set exactly, provided
// the box test is conservative with respect to the FP32 rounding of stage A.  That slack is applied
// per ray in the kernel (it depends on the ray's origin); here the only per-patch input is the
// conditioning kappa = max_i |row_i(Binv)|_1 * altitude_i, which scales the barycentric rounding
// error into a distance: the in-plane error of the inside test is at most kappa * 2^-19 * (|s|_1 + 2 extent)
// for a ray starting at s.  The kernel's per-ray slack covers kappa <= 16.  Patches with
// 16 < kappa <= 65536 ("wide": underlying plane seen at a grazing angle from the origin) stay in the
// hierarchy with their boxes grown on the host by the remaining (kappa - 16) share, evaluated for the
// largest ray origin the hierarchy vouches for (origin_limit); a ray starting farther out tests them
// directly.  Patches with a larger or non-finite kappa go to the "always" segment and are tested for
// every ray, as the reference does.
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <limits>
#include <vector>
#include "bz_bvh.h"

#ifndef BZ_BVH_LEAF_MAX
#define BZ_BVH_LEAF_MAX 2     // = BZ_LEAF_MAX of csrc/bz_device.h (bz_capi.cu checks it)
#endif

namespace {

struct Box {
  float lo[3], hi[3];
  void reset() { for(int k = 0; k < 3; ++k) { lo[k] = std::numeric_limits<float>::max(); hi[k] = -std::numeric_limits<float>::max(); } }
  void grow(const float *p) { for(int k = 0; k < 3; ++k) { lo[k] = std::min(lo[k], p[k]); hi[k] = std::max(hi[k], p[k]); } }
  void grow(const Box &b) { for(int k = 0; k < 3; ++k) { lo[k] = std::min(lo[k], b.lo[k]); hi[k] = std::max(hi[k], b.hi[k]); } }
  double area() const {
    double dx = double(hi[0]) - lo[0], dy = double(hi[1]) - lo[1], dz = double(hi[2]) - lo[2];
    if(dx < 0.0 || dy < 0.0 || dz < 0.0) { return 0.0; }
    return 2.0 * (dx * dy + dy * dz + dz * dx);
  }
};

struct Binary {
  Box box;
  int left, right;          // -1 for leaves
  uint32_t first, count;    // leaf range in `order`
};

constexpr int cBins = 16;
constexpr uint32_t cLeafMax = BZ_BVH_LEAF_MAX;   // <= 7 (3 bits in the leaf code)
constexpr double cKappaLimit = 16.0;
constexpr double cKappaWideLimit = 65536.0;
constexpr double cSlackUnit = 1.9073486328125e-06;     // 2^-19; the kernel's per-ray slack is cKappaLimit * cSlackUnit * (|s|_1 + 2 extent)
constexpr double cOriginLimitFactor = 64.0;

struct Builder {
  std::vector<Box> boxes;            // per primitive (index into prims)
  std::vector<float> centroid;       // 3 per primitive
  std::vector<uint32_t> order;       // permutation being partitioned
  std::vector<Binary> nodes;

  int build(uint32_t begin, uint32_t end, int depth) {
    Binary node;
    node.box.reset();
    Box cbox;
    cbox.reset();
    for(uint32_t i = begin; i < end; ++i) {
      node.box.grow(boxes[order[i]]);
      cbox.grow(&centroid[3u * order[i]]);
    }
    node.left = node.right = -1;
    node.first = begin;
    node.count = end - begin;
    int const index = static_cast<int>(nodes.size());
    nodes.push_back(node);
    if(end - begin <= cLeafMax) { return index; }

    uint32_t mid = begin;
    bool split = false;
    if(depth < 40) {
      double best = std::numeric_limits<double>::max();
      int bestAxis = -1, bestBin = -1;
      for(int axis = 0; axis < 3; ++axis) {
        double const lo = cbox.lo[axis], extent = double(cbox.hi[axis]) - lo;
        if(!(extent > 0.0)) { continue; }
        Box bin[cBins];
        uint32_t cnt[cBins] = {};
        for(auto &b : bin) { b.reset(); }
        for(uint32_t i = begin; i < end; ++i) {
          int k = std::min(cBins - 1, static_cast<int>((centroid[3u * order[i] + axis] - lo) / extent * cBins));
          bin[k].grow(boxes[order[i]]);
          ++cnt[k];
        }
        double rightArea[cBins];
        Box acc;
        acc.reset();
        uint32_t n = 0u;
        uint32_t rightCount[cBins];
        for(int k = cBins - 1; k > 0; --k) {
          acc.grow(bin[k]);
          n += cnt[k];
          rightArea[k] = acc.area();
          rightCount[k] = n;
        }
        acc.reset();
        n = 0u;
        for(int k = 0; k < cBins - 1; ++k) {
          acc.grow(bin[k]);
          n += cnt[k];
          if(n == 0u || rightCount[k + 1] == 0u) { continue; }
          double cost = acc.area() * n + rightArea[k + 1] * rightCount[k + 1];
          if(cost < best) { best = cost; bestAxis = axis; bestBin = k; }
        }
      }
      if(bestAxis >= 0) {
        double const lo = cbox.lo[bestAxis], extent = double(cbox.hi[bestAxis]) - lo;
        auto middle = std::partition(order.begin() + begin, order.begin() + end, [&](uint32_t p) {
          int k = std::min(cBins - 1, static_cast<int>((centroid[3u * p + bestAxis] - lo) / extent * cBins));
          return k <= bestBin;
        });
        mid = static_cast<uint32_t>(middle - order.begin());
        split = (mid > begin && mid < end);
      }
    }
    if(!split) {
      mid = begin + (end - begin) / 2u;   // coincident centroids or depth cap: split the index range
    }
    int const left = build(begin, mid, depth + 1);
    int const right = build(mid, end, depth + 1);
    nodes[index].left = left;
    nodes[index].right = right;
    nodes[index].count = 0u;
    return index;
  }
};

}



void bz_build_bvh(const bz_patch *patches, uint32_t count, bz_bvh &out) {
  out.nodes.clear();
  out.order.clear();
  out.always.clear();
  out.extent = 0.0f;
  out.depth = 0u;
  out.wide.clear();
  out.origin_limit = 0.0f;
  out.packet_nodes.clear();
  out.packet_depth = 0u;
  Builder b;
  std::vector<uint32_t> prims;     // primitive -> patch
  std::vector<double> kappas(count, std::numeric_limits<double>::infinity());
  for(uint32_t i = 0u; i < count; ++i) {
    const bz_patch &p = patches[i];
    const float *v[3] = { p.control[0], p.control[1], p.control[2] };
    bool finite = true;
    for(int a = 0; a < 3; ++a) {
      for(int k = 0; k < 3; ++k) { finite = finite && std::isfinite(v[a][k]); }
    }
    if(!finite) { continue; }
    double e[3][3];   // e[i] = edge opposite vertex i
    for(int k = 0; k < 3; ++k) {
      e[0][k] = double(v[2][k]) - v[1][k];
      e[1][k] = double(v[0][k]) - v[2][k];
      e[2][k] = double(v[1][k]) - v[0][k];
    }
    double cx = e[2][1] * -e[1][2] - e[2][2] * -e[1][1];
    double cy = e[2][2] * -e[1][0] - e[2][0] * -e[1][2];
    double cz = e[2][0] * -e[1][1] - e[2][1] * -e[1][0];
    double const twiceArea = std::sqrt(cx * cx + cy * cy + cz * cz);
    double kappa = 0.0;
    for(int i3 = 0; i3 < 3; ++i3) {
      double const edge = std::sqrt(e[i3][0] * e[i3][0] + e[i3][1] * e[i3][1] + e[i3][2] * e[i3][2]);
      double const altitude = twiceArea / edge;
      double const row = std::fabs(double(p.bary_inv[i3][0])) + std::fabs(double(p.bary_inv[i3][1])) + std::fabs(double(p.bary_inv[i3][2]));
      kappa = std::max(kappa, row * altitude);
      if(!(row * altitude == row * altitude)) { kappa = std::numeric_limits<double>::infinity(); }
    }
    kappas[i] = kappa;
    for(int a = 0; a < 3; ++a) {
      for(int k = 0; k < 3; ++k) { out.extent = std::max(out.extent, std::fabs(v[a][k])); }
    }
  }
  out.origin_limit = static_cast<float>(cOriginLimitFactor * out.extent);
  for(uint32_t i = 0u; i < count; ++i) {
    const bz_patch &p = patches[i];
    const float *v[3] = { p.control[0], p.control[1], p.control[2] };
    if(!(kappas[i] <= cKappaWideLimit)) {
      out.always.push_back(i);
      continue;
    }
    Box box;
    box.reset();
    for(int a = 0; a < 3; ++a) { box.grow(v[a]); }
    if(kappas[i] > cKappaLimit) {
      // 1.5: the sum below is rounded to float twice (grow, then the kernel's own arithmetic)
      float const grow = static_cast<float>(1.5 * (kappas[i] - cKappaLimit) * cSlackUnit * (double(out.origin_limit) + 2.0 * out.extent));
      for(int k = 0; k < 3; ++k) { box.lo[k] -= grow; box.hi[k] += grow; }
      out.wide.push_back(i);
    }
    b.boxes.push_back(box);
    for(int k = 0; k < 3; ++k) { b.centroid.push_back((v[0][k] + v[1][k] + v[2][k]) / 3.0f); }
    prims.push_back(i);
  }
  uint32_t const n = static_cast<uint32_t>(prims.size());
  if(n == 0u) { return; }
  b.order.resize(n);
  for(uint32_t i = 0u; i < n; ++i) { b.order[i] = i; }
  b.nodes.reserve(2u * n);
  int const root = b.build(0u, n, 0);
  out.order.resize(n);
  for(uint32_t i = 0u; i < n; ++i) { out.order[i] = prims[b.order[i]]; }

  // flatten in depth-first order with skip links: node i is followed by its first child (inner node) or by the next
  // subtree (leaf); link = index of the first node after the subtree of i, taken when the box is missed
  struct Frame { int binary; size_t flat; uint32_t depth; int stage; };
  std::vector<Frame> todo;
  todo.push_back({ root, 0u, 1u, 0 });
  while(!todo.empty()) {
    Frame &job = todo.back();
    Binary const &bin = b.nodes[job.binary];
    if(job.stage == 0) {
      job.flat = out.nodes.size() / 8u;
      out.nodes.resize(out.nodes.size() + 8u, 0.0f);
      float *p = out.nodes.data() + job.flat * 8u;
      p[0] = bin.box.lo[0]; p[1] = bin.box.lo[1]; p[2] = bin.box.lo[2];
      p[4] = bin.box.hi[0]; p[5] = bin.box.hi[1]; p[6] = bin.box.hi[2];
      out.depth = std::max(out.depth, job.depth);
      if(bin.left < 0) {
        uint32_t const code = (bin.first << 3u) | bin.count;      // never 0: count >= 1
        std::memcpy(p + 7, &code, sizeof(code));
        job.stage = 3;
      }
      else {
        job.stage = 1;
        int const child = bin.left;
        uint32_t const depth = job.depth + 1u;
        todo.push_back({ child, 0u, depth, 0 });                 // invalidates `job`
      }
    }
    else if(job.stage == 1) {
      job.stage = 3;
      int const child = bin.right;
      uint32_t const depth = job.depth + 1u;
      todo.push_back({ child, 0u, depth, 0 });
    }
    else {
      uint32_t link = static_cast<uint32_t>(out.nodes.size() / 8u);      // patched to the end marker below if nothing follows
      std::memcpy(out.nodes.data() + job.flat * 8u + 3u, &link, sizeof(link));
      todo.pop_back();
    }
  }
  // the same tree collapsed to 32 children per node: a node's child set starts as its two binary children and the
  // inner member with the largest box is replaced by its own two children until 32 are reached or only leaves remain
  {
    struct Wide { int binary; uint32_t depth; };
    std::vector<Wide> queue;
    queue.push_back({ root, 1u });
    for(size_t at = 0u; at < queue.size(); ++at) {
      Wide const job = queue[at];
      out.packet_depth = std::max(out.packet_depth, job.depth);
      std::vector<int> members;
      if(b.nodes[job.binary].left < 0) { members.push_back(job.binary); }
      else { members.push_back(b.nodes[job.binary].left); members.push_back(b.nodes[job.binary].right); }
      while(members.size() < 32u) {
        int pick = -1;
        double area = -1.0;
        for(size_t m = 0u; m < members.size(); ++m) {
          Binary const &cand = b.nodes[members[m]];
          if(cand.left >= 0 && cand.box.area() > area) { area = cand.box.area(); pick = static_cast<int>(m); }
        }
        if(pick < 0) { break; }
        Binary const &open = b.nodes[members[pick]];
        members[pick] = open.left;
        members.push_back(open.right);
      }
      size_t const base = at * 224u;
      if(out.packet_nodes.size() < base + 224u) { out.packet_nodes.resize(base + 224u, 0.0f); }
      for(uint32_t lane = 0u; lane < 32u; ++lane) {
        float *p = out.packet_nodes.data() + base;
        uint32_t code = 0u;
        if(lane < members.size()) {
          Binary const &child = b.nodes[members[lane]];
          for(int k = 0; k < 3; ++k) { p[32 * k + lane] = child.box.lo[k]; p[32 * (3 + k) + lane] = child.box.hi[k]; }
          if(child.left < 0) { code = (child.first << 3u) | child.count; }
          else {
            code = 0x80000000u | static_cast<uint32_t>(queue.size());
            queue.push_back({ members[lane], job.depth + 1u });
            out.packet_nodes.resize(queue.size() * 224u, 0.0f);
            p = out.packet_nodes.data() + base;
          }
        }
        else {
          for(int k = 0; k < 3; ++k) { p[32 * k + lane] = 3.0e38f; p[32 * (3 + k) + lane] = -3.0e38f; }
        }
        std::memcpy(p + 192 + lane, &code, sizeof(code));
      }
    }
  }
  uint32_t const total = static_cast<uint32_t>(out.nodes.size() / 8u), endMarker = 0xffffffffu;
  for(uint32_t i = 0u; i < total; ++i) {
    uint32_t link;
    std::memcpy(&link, out.nodes.data() + size_t(i) * 8u + 3u, sizeof(link));
    if(link >= total) { std::memcpy(out.nodes.data() + size_t(i) * 8u + 3u, &endMarker, sizeof(endMarker)); }
  }
}
