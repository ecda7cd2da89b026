// bz_hemisphere.h -- host C++ face of the emitter sampler: the reference's UniformHemisphere (reference/hostUtil.h:8-28)
// with the same constructor argument and getPatchCount(); the sampling itself runs on the device (bz_emit_rays,
// csrc/bz_illum.cu), a batch of rays at a time, with the counter-based generator of include/bz_capi.h in place of the
// reference's std::ranlux24_base member (statistical, not bit, parity with the reference's stream; bit parity with
// oracle/bz_oracle_illum.c).  Header-only; link against libbz_b200.so.
#ifndef BZ_HEMISPHERE_H
#define BZ_HEMISPHERE_H

#include <cstdint>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>
#include "bz_capi.h"

namespace bz {

class UniformHemisphere final {
public:
  // UniformHemisphere(uint32_t const aBelts), reference/hostUtil.cpp:3-14
  explicit UniformHemisphere(uint32_t const aBelts) : mBeltCount(aBelts), mWidths(aBelts), mFirsts(aBelts) {
    check(bz_hemisphere_patches(aBelts, &mPatchCount, mWidths.data(), mFirsts.data()));
  }

  uint32_t getPatchCount() const { return mPatchCount; }                       // reference/hostUtil.h:23
  uint32_t getBeltCount() const { return mBeltCount; }
  // mBelts[i] of the reference: (patch width of belt i, patches before belt i)
  std::pair<float, uint32_t> getBelt(uint32_t const aIndex) const { return std::make_pair(mWidths.at(aIndex), mFirsts.at(aIndex)); }

  // getRandom() (reference/hostUtil.cpp:16-29) for rays [aFirst, aFirst + aCount) of aEmitter's emission order, into DEVICE
  // arrays (3 floats per ray; aPatch may be null); aEmitter.belts must be this object's belt count
  void getRandom(bz_emitter const &aEmitter, uint64_t const aFirst, uint64_t const aCount, float *aDeviceStart, float *aDeviceDirection,
                 uint32_t *aDevicePatch, void *aStream = nullptr) const {
    if(aEmitter.belts != mBeltCount) { throw std::invalid_argument("emitter and hemisphere disagree on the belt count"); }
    check(bz_emit_rays(&aEmitter, aFirst, aCount, aDeviceStart, aDeviceDirection, aDevicePatch, aStream));
  }

private:
  uint32_t              mBeltCount;
  uint32_t              mPatchCount = 0u;
  std::vector<float>    mWidths;
  std::vector<uint32_t> mFirsts;

  static void check(int const aCode) {
    if(aCode != BZ_OK) { throw std::runtime_error("libbz_b200 error " + std::to_string(aCode) + ": " + bz_last_error()); }
  }
};

}

#endif
