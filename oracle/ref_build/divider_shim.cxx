// TEST INFRASTRUCTURE -- extern "C" shim around the UNMODIFIED reference class divalgo::DividerAlgo
// (/root/reference/caffe/larcv/app/HiResDivider/DividerAlgo.{h,cxx}, WireInfoData.h), compiled where the sources lie by
// oracle/ref_build/Makefile into oracle/_ref/libdivider_ref.so.  Used only by tests (to pin
// ubresnet_b200/deploy/crop_geometry.py) -- never by the product path.
#include <vector>
#include "DividerAlgo.h"

extern "C" {

// returns 0 and fills out[6] = {y_lo, y_hi, u_lo, u_hi, v_lo, v_hi} (first / last wire of each list; -1,-1 when empty);
// returns 1 when the reference throws (box corners that "don't make sense")
int ref_getregionwires(float z1, float y1, float z2, float y2, int* out) {
  divalgo::DividerAlgo algo;
  std::vector<int> yw, uw, vw;
  try {
    algo.getregionwires(z1, y1, z2, y2, yw, uw, vw);
  } catch (...) {
    return 1;
  }
  const std::vector<int>* v[3] = {&yw, &uw, &vw};
  for (int i = 0; i < 3; ++i) {
    out[2 * i] = v[i]->empty() ? -1 : v[i]->front();
    out[2 * i + 1] = v[i]->empty() ? -1 : v[i]->back();
  }
  return 0;
}

// the handful of detector constants the algorithm reads from the wire table (WireInfoData.h), for the pin test
void ref_wire_constants(float* out) {
  using namespace divalgo;
  int k = 0;
  out[k++] = uplaneStart[0][1];    out[k++] = uplaneStart[0][2];
  out[k++] = uplaneStart[2399][1]; out[k++] = uplaneStart[2399][2];
  out[k++] = uplaneStart[100][1];  out[k++] = uplaneStart[101][1];  out[k++] = uplaneStart[100][2];
  out[k++] = uplaneStart[1000][1]; out[k++] = uplaneStart[1000][2]; out[k++] = uplaneStart[1001][2];
  out[k++] = vplaneStart[2300][1]; out[k++] = vplaneStart[2301][1]; out[k++] = vplaneStart[2300][2];
  out[k++] = vplaneStart[100][1];  out[k++] = vplaneStart[100][2];  out[k++] = vplaneStart[101][2];
  out[k++] = yplaneStart[0][2];
}

}
