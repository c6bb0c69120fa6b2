// Detection -> optimiser-input builder (frame.cu): device-side parameter blocks and launchers.
#pragma once
#include <cuda_runtime.h>

#include <cstddef>
#include <cstdint>

namespace dsp {

constexpr int FB_MAXB = 32;          // detections per count / emit launch (one bit each)
constexpr int FB_MAX_DETS = 256;     // detections per frame
constexpr int FB_MAX_MASKS = 256;    // 2-D instance masks per frame
constexpr int FB_INFO = 8;           // int32 per detection in the info output
enum { FBI_IN_BOX = 0, FBI_M = 1, FBI_NFOV = 2, FBI_MASK = 3, FBI_HAS_RAYS = 4, FBI_NBG = 5, FBI_HITS = 6 };

struct FrameParamsDev {
    double Kd[9], invKd[9];          // monocular path: FP64 intrinsics (mono_sequence.py:130-131)
    int img_w, img_h, max_pts, min_mask_area, alpha, max_bg;
    float K[9], invK[9], Rcv[9], tcv[3];
    double k1, k2;
};
struct FrameBoxDev {
    float Rov[9], tov[3], lo[3], hi[3], half[3];
};
struct FrameBoxes {
    FrameBoxDev b[FB_MAXB];
    int n;
};
struct FrontBits {
    uint32_t w[FB_MAX_DETS / 32];
};

size_t frame_scratch_bytes(int64_t n_velo, int B, int n_masks, int max_bg);
int launch_frame_kitti(const FrameParamsDev& prm, const float* velo, int64_t n_velo, const FrameBoxDev* boxes,
                       const int* is_front, int B, const uint8_t* masks, int n_masks, const float* bboxes,
                       int* pts_off, int* ray_off, int* n_fg, float* pts, float* rays, float* depth, int* info,
                       void* scratch, cudaStream_t st);
int launch_frame_mono(const FrameParamsDev& prm, const uint8_t* masks, int n_masks, const float* bboxes, float* rays,
                      int* info, void* scratch, cudaStream_t st);

}  // namespace dsp
