/*
 * mg_frames.cu -- the preprocess half of the step path applied to STORED full-resolution frames
 * (demo trajectories on disk): ResizeDictObservation's cv2.resize(INTER_AREA) at exactly 4x
 * (magical/benchmarks/__init__.py:150-190) and FlattenFrameStack / EagerDictFrameStack / ChannelsFirst
 * (magical/benchmarks/__init__.py:51-147, 193-216) as preprocess_demos_with_wrapper replays them over a
 * trajectory (magical/saved_trajectories.py:87-149).  Handle-less entry points of the C ABI
 * (include/mg_b200.h: mg_frames_downsample, mg_frames_stack); both kernels are pure HBM streams.
 */
#include <cuda_runtime.h>
#include <stdint.h>
#include <string>
#include "mg_kernels.h"

#define FR_RES MG_RES
#define FR_LO MG_LORES

int mg_fail_global(int code, const std::string &msg);   /* mg_api.cu: message for mg_last_error(NULL) */

/* round-half-even(sum / 16) of three channel sums packed for one output pixel */
__device__ __forceinline__ uint32_t rhe16(uint32_t s)
{
    uint32_t q = s >> 4, r = s & 15;
    return q + ((r > 8 || (r == 8 && (q & 1))) ? 1u : 0u);
}

/* One thread = 4 neighbouring output pixels of one output row: per input row it loads 16 pixels = 48 bytes
 * as three 16-byte words (a warp reads 24 x 48 contiguous bytes of each of the 4 rows), sums the channels with
 * byte-select dot products and stores 12 output bytes as three words. */
__global__ void __launch_bounds__(256) mg_frames_downsample_kernel(const uint8_t *__restrict__ in, uint8_t *__restrict__ out, long long n_items)
{
    const long long item = (long long)blockIdx.x * blockDim.x + threadIdx.x;     /* (frame, out row, quad) */
    if (item >= n_items) return;
    const int quad = (int)(item % (FR_LO / 4));
    const long long row_id = item / (FR_LO / 4);
    const int oy = (int)(row_id % FR_LO);
    const long long frame = row_id / FR_LO;
    const uint4 *src = (const uint4 *)(in + ((size_t)frame * FR_RES + 4 * oy) * FR_RES * 3) + quad * 3;
    uint32_t sum[4][3];
#pragma unroll
    for (int p = 0; p < 4; p++) sum[p][0] = sum[p][1] = sum[p][2] = 0;
#pragma unroll
    for (int r = 0; r < 4; r++) {
        const uint4 *rp = src + (size_t)r * (FR_RES * 3 / 16);
        uint4 q0 = __ldg(rp), q1 = __ldg(rp + 1), q2 = __ldg(rp + 2);
        const uint32_t w[12] = { q0.x, q0.y, q0.z, q0.w, q1.x, q1.y, q1.z, q1.w, q2.x, q2.y, q2.z, q2.w };
#pragma unroll
        for (int p = 0; p < 4; p++) {
            /* 4 input pixels = 12 bytes R G B R | G B R G | B R G B */
            const uint32_t a = w[3 * p], b = w[3 * p + 1], c = w[3 * p + 2];
            sum[p][0] = __dp4a(a, 0x01000001u, __dp4a(b, 0x00010000u, __dp4a(c, 0x00000100u, sum[p][0])));
            sum[p][1] = __dp4a(a, 0x00000100u, __dp4a(b, 0x01000001u, __dp4a(c, 0x00010000u, sum[p][1])));
            sum[p][2] = __dp4a(a, 0x00010000u, __dp4a(b, 0x00000100u, __dp4a(c, 0x01000001u, sum[p][2])));
        }
    }
    uint32_t o[12];
#pragma unroll
    for (int p = 0; p < 4; p++)
#pragma unroll
        for (int ch = 0; ch < 3; ch++) o[3 * p + ch] = rhe16(sum[p][ch]);
    uint32_t *dst = (uint32_t *)(out + ((size_t)frame * FR_LO + oy) * FR_LO * 3) + quad * 3;
#pragma unroll
    for (int k = 0; k < 3; k++) dst[k] = o[4 * k] | o[4 * k + 1] << 8 | o[4 * k + 2] << 16 | o[4 * k + 3] << 24;
}

struct StackSpec { int n_groups, chw; int view[MG_MAX_GROUPS], age[MG_MAX_GROUPS]; };

/* out[t] = channel-concatenation of the frames (view g, max(t - age g, 0)): the deque a stacking wrapper holds
 * after `t` steps, its reset having pushed frame 0 `depth` times.  One thread per output pixel. */
__global__ void __launch_bounds__(256) mg_frames_stack_kernel(const uint8_t *__restrict__ lo0, const uint8_t *__restrict__ lo1,
                                                               StackSpec sp, uint8_t *__restrict__ out, long long n_pix)
{
    const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= n_pix) return;
    const int pix = (int)(idx % (FR_LO * FR_LO));
    const long long t = idx / (FR_LO * FR_LO);
    const int C = 3 * sp.n_groups;
    uint8_t *o = out + (size_t)t * FR_LO * FR_LO * C;
    for (int g = 0; g < sp.n_groups; g++) {
        const long long ts = t - sp.age[g] < 0 ? 0 : t - sp.age[g];
        const uint8_t *s = (sp.view[g] == MG_VIEW_ALLO ? lo0 : lo1) + ((size_t)ts * FR_LO * FR_LO + pix) * 3;
        for (int c = 0; c < 3; c++) {
            const int ch = 3 * g + c;
            if (sp.chw) o[(size_t)ch * FR_LO * FR_LO + pix] = s[c];
            else o[(size_t)pix * C + ch] = s[c];
        }
    }
}

extern "C" int mg_frames_downsample(int device, const uint8_t *frames_dev, int64_t n_frames, uint8_t *out_dev, void *stream)
{
    if (n_frames < 0 || (n_frames > 0 && (!frames_dev || !out_dev))) return mg_fail_global(MG_ERR_ARG, "mg_frames_downsample: bad argument");
    if (((uintptr_t)frames_dev & 15) || ((uintptr_t)out_dev & 3)) return mg_fail_global(MG_ERR_ARG, "mg_frames_downsample: frames must be 16-byte aligned");
    if (n_frames == 0) return MG_OK;
    if (cudaSetDevice(device) != cudaSuccess) return mg_fail_global(MG_ERR_NODEVICE, "mg_frames_downsample: no such CUDA device (there is no CPU path)");
    const long long items = (long long)n_frames * FR_LO * (FR_LO / 4);
    mg_frames_downsample_kernel<<<(unsigned)((items + 255) / 256), 256, 0, (cudaStream_t)stream>>>(frames_dev, out_dev, items);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return mg_fail_global(MG_ERR_CUDA, std::string("mg_frames_downsample: ") + cudaGetErrorString(e));
    return MG_OK;
}

extern "C" int mg_frames_stack(int device, const uint8_t *allo_lo_dev, const uint8_t *ego_lo_dev, int64_t n_frames, int n_groups,
                               const int32_t *group_view, const int32_t *group_age, int chw, uint8_t *out_dev, void *stream)
{
    if (n_frames < 0 || n_groups <= 0 || n_groups > MG_MAX_GROUPS || !group_view || !group_age || (n_frames > 0 && !out_dev))
        return mg_fail_global(MG_ERR_ARG, "mg_frames_stack: bad argument");
    StackSpec sp;
    sp.n_groups = n_groups; sp.chw = chw;
    for (int g = 0; g < n_groups; g++) {
        sp.view[g] = group_view[g]; sp.age[g] = group_age[g];
        if (sp.age[g] < 0 || (sp.view[g] != MG_VIEW_ALLO && sp.view[g] != MG_VIEW_EGO)) return mg_fail_global(MG_ERR_ARG, "mg_frames_stack: bad group");
        if (n_frames > 0 && !(sp.view[g] == MG_VIEW_ALLO ? allo_lo_dev : ego_lo_dev)) return mg_fail_global(MG_ERR_ARG, "mg_frames_stack: a group names a view without frames");
    }
    if (n_frames == 0) return MG_OK;
    if (cudaSetDevice(device) != cudaSuccess) return mg_fail_global(MG_ERR_NODEVICE, "mg_frames_stack: no such CUDA device (there is no CPU path)");
    const long long n_pix = (long long)n_frames * FR_LO * FR_LO;
    mg_frames_stack_kernel<<<(unsigned)((n_pix + 255) / 256), 256, 0, (cudaStream_t)stream>>>(allo_lo_dev, ego_lo_dev, sp, out_dev, n_pix);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return mg_fail_global(MG_ERR_CUDA, std::string("mg_frames_stack: ") + cudaGetErrorString(e));
    return MG_OK;
}
