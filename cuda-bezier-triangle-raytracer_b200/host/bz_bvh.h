// bz_bvh.h -- exact cull hierarchy over the flat underlying triangles of the patches (host build).
#ifndef BZ_BVH_H
#define BZ_BVH_H

#include <cstdint>
#include <vector>
#include "bz_capi.h"

struct bz_bvh {
  std::vector<float>    nodes;    // 8 floats per node in depth-first order: lo.xyz, skip link, hi.xyz, leaf code (csrc/bz_device.h)
  std::vector<uint32_t> order;    // leaf slot -> patch index
  std::vector<uint32_t> wide;     // ill-conditioned patches kept in the hierarchy with grown boxes; a ray whose origin has
                                  // |s|_1 > origin_limit tests them directly as well
  std::vector<uint32_t> always;   // patches outside the hierarchy, tested for every ray
  float                 extent;   // max |coordinate| of the flat triangles
  float                 origin_limit;
  uint32_t              depth;    // levels of the binary hierarchy
  // the same hierarchy collapsed to 32 children per node, for the warp-cooperative packet walk (one lane per child):
  // 7 x 32 words per node: lo.x[32] lo.y[32] lo.z[32] hi.x[32] hi.y[32] hi.z[32] code[32]
  // code: 0 = empty child, bit 31 set = inner child (low bits: packet node index), else first_slot << 3 | count
  std::vector<float>    packet_nodes;
  uint32_t              packet_depth;
};

void bz_build_bvh(const bz_patch *patches, uint32_t count, bz_bvh &out);

#endif
