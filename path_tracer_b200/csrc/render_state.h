// render_state.h — what other translation units may ask of the frame state owned by render.cu.
#pragma once
#include <cuda_runtime.h>

#include <cstdint>

struct ptb_context;
namespace ptb {
// current accumulation buffer (bound or own) and frame size; false when ptb_frame_begin has not been called
bool render_state_frame(ptb_context* ctx, float4** accum, uint32_t* W, uint32_t* H);
}  // namespace ptb
