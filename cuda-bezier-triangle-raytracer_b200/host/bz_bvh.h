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
};

void bz_build_bvh(const bz_patch *patches, uint32_t count, bz_bvh &out);

#endif
