// NVTX range named after the C-ABI entry it sits in (what nsys shows where the reference has CV_INSTRUMENT_REGION)
#pragma once
#include <nvtx3/nvToolsExt.h>
namespace b200 {
struct NvtxRange {
    explicit NvtxRange(const char* name) { nvtxRangePushA(name); }
    ~NvtxRange() { nvtxRangePop(); }
};
}
#define B200_TRACE() ::b200::NvtxRange nvtx_range_(__func__)
