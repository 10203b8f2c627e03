// renderer_internal.hpp -- what the other translation units of the library may see of a ptb_renderer (renderer.cu).
#pragma once
#include <cuda_runtime.h>

#include "../cuda/render.cuh"

struct ptb_renderer;
const ptb::FrameDev& ptb_internal_frame(const ptb_renderer* r);
int ptb_internal_device(const ptb_renderer* r);
cudaStream_t ptb_internal_stream(const ptb_renderer* r);
void ptb_internal_count_launches(ptb_renderer* r, int n);
int ptb_internal_num_sms(const ptb_renderer* r);
