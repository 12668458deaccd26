// ref_bev_host.cu — builds the reference's OWN rotated-IoU math for the host.
// TEST INFRASTRUCTURE ONLY (see oracle/README.md).
//
// bev_iou/bev_iou_g.cu declares box_overlap()/iou_bev() as __device__ only.
// Re-defining __device__ as "__host__ __device__" while including the file where
// it lies under /root/reference lets nvcc emit host versions of exactly that
// source, which the two thin extern "C" shims below export.  No reference source
// is copied; the include path is passed by oracle/Makefile (-DREF_BEV_IOU_G_CU=...).
#include <cuda_runtime.h>
#undef __device__
#define __device__ __host__ __location__(device)
#include REF_BEV_IOU_G_CU

extern "C" float ref_host_box_overlap(const float *a, const float *b) { return box_overlap(a, b); }
extern "C" float ref_host_iou_bev(const float *a, const float *b) { return iou_bev(a, b); }
