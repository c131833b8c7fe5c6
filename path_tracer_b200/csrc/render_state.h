// render_state.h — what other translation units may ask of the frame state owned by render.cu.
#pragma once
#include <cuda_runtime.h>

#include <cstdint>

struct ptb_context;
namespace ptb {
// current accumulation buffer (bound or own) and frame size; false when ptb_frame_begin has not been called
bool render_state_frame(ptb_context* ctx, float4** accum, uint32_t* W, uint32_t* H);
// first-hit position (xyz, t) and packed id history ((previous << 16) | current, main.rs:206) of the current frame
bool render_state_aux(ptb_context* ctx, float4** position, uint32_t** id);
// accumulation texture of the temporal pass (temporal.cu); false before the first ptb_state_* call
bool temporal_state_image(ptb_context* ctx, const float4** history, uint32_t* W, uint32_t* H);
}  // namespace ptb
