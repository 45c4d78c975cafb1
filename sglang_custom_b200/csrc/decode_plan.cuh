// Layout of the device-side decode plan (shared by decode_plan.cu and decode.cu).
#pragma once
#include <stdint.h>

namespace b200attn {

constexpr int kUnit = 16;  // tokens per scheduling unit

// int32 header
constexpr int kPlanTotalUnits = 0;   // sum_i ceil(seq_len_i / kUnit)
constexpr int kPlanUnitsPerCta = 1;  // W: CTA c owns units [c*W, (c+1)*W)
constexpr int kPlanBs = 2;
constexpr int kPlanNumCtas = 3;
constexpr int kPlanWindow = 4;       // sliding_window_size the units were counted for (-1: whole sequence)
constexpr int kPlanHeader = 8;
// then: cum[bs + 1] (exclusive prefix sum of units), first_req[n_ctas]

}  // namespace b200attn
