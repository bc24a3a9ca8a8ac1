// LiDAR condition pipeline on the device.
//
// The reference builds every condition frame on the host: NumPy concatenation of the +-delta_frames
// background clouds and of every active actor's clouds, a float64 homogeneous transform of the actor
// points by ego_pose @ box pose, a host->device copy, the splat, a device->host copy, uint8 quantisation
// and two PNG files that the diffusion runner reads back
// (/root/reference/street_gaussian/pointcloud_processor/base_processor.py:32-63,
//  /root/reference/street_gaussian/pointcloud_processor/waymo_processor.py:179-242,
//  /root/reference/data_processor/utils/render_utils.py:83-126).
// Here the clouds live on the device once; per frame ONE gather kernel writes the point set of the
// frame (positions, colours, isotropic splat scales) straight into the arrays the projection kernel
// reads, and one pack kernel turns the compositor's output into the uint8 [H,W,4] image the PNG pair
// holds (and / or the float [4,H,W] tensor the diffusion model consumes).
#include "common.cuh"

namespace {

constexpr int LIDAR_MAX_SEGS = 512;

// one contiguous run of source points that lands at [dst0, dst0 + count) of the frame's point set
struct LidarSeg {
    int64_t dst0;
    int64_t count;
    const void* src;      // rows of 6 values: x y z r g b (float32 for world-space clouds, float64 for actors)
    int32_t is_f64;
    int32_t has_pose;
    double pose[12];      // rows 0..2 of the 4x4 object->world matrix (float64, as the reference multiplies)
};

__global__ void __launch_bounds__(256)
lidar_gather_kernel(int64_t n, int n_segs, const LidarSeg* __restrict__ segs, const float* __restrict__ w2c_row2,
                    float scale, float fx, float min_hw, int ndc_scale, float* __restrict__ means,
                    float* __restrict__ rgb, float* __restrict__ scales) {
    __shared__ int64_t s_dst0[LIDAR_MAX_SEGS + 1];
    for (int i = threadIdx.x; i <= n_segs; i += blockDim.x) s_dst0[i] = i < n_segs ? segs[i].dst0 : n;
    __syncthreads();
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    int lo = 0, hi = n_segs;  // s_dst0[lo] <= i < s_dst0[hi]
    while (hi - lo > 1) {
        const int mid = (lo + hi) >> 1;
        if (s_dst0[mid] <= i) lo = mid; else hi = mid;
    }
    const LidarSeg& sg = segs[lo];
    const int64_t r = i - sg.dst0;
    float x, y, z, cr, cg, cb;
    if (sg.is_f64) {
        const double* p = reinterpret_cast<const double*>(sg.src) + r * 6;
        double px = p[0], py = p[1], pz = p[2];
        if (sg.has_pose) {
            // (xyz_homo @ pose.T)[:3] in float64, rounded to float32 afterwards (torch .float())
            const double qx = px * sg.pose[0] + py * sg.pose[1] + pz * sg.pose[2] + sg.pose[3];
            const double qy = px * sg.pose[4] + py * sg.pose[5] + pz * sg.pose[6] + sg.pose[7];
            const double qz = px * sg.pose[8] + py * sg.pose[9] + pz * sg.pose[10] + sg.pose[11];
            px = qx; py = qy; pz = qz;
        }
        x = (float)px; y = (float)py; z = (float)pz;
        cr = (float)p[3]; cg = (float)p[4]; cb = (float)p[5];
    } else {
        const float* p = reinterpret_cast<const float*>(sg.src) + r * 6;
        x = p[0]; y = p[1]; z = p[2];
        cr = p[3]; cg = p[4]; cb = p[5];
    }
    means[3 * i + 0] = x; means[3 * i + 1] = y; means[3 * i + 2] = z;
    rgb[3 * i + 0] = cr; rgb[3 * i + 1] = cg; rgb[3 * i + 2] = cb;
    float s = scale;
    if (ndc_scale) {
        // scales * z / fx * 0.5 * min(H, W)   (render_utils.py:116-122), z = view-space depth
        const float vz = x * w2c_row2[0] + y * w2c_row2[1] + z * w2c_row2[2] + w2c_row2[3];
        s = s * vz / fx;
        s = s * 0.5f * min_hw;
    }
    scales[3 * i + 0] = s; scales[3 * i + 1] = s; scales[3 * i + 2] = s;
}

// compositor output (rgb[H,W,3], alpha[H,W,1]) -> uint8 [H,W,4] exactly as (x * 255).astype(np.uint8)
// (waymo_processor.py:237-238) and / or float [4,H,W] in [0,1] from the QUANTISED values (what PIL +
// ToTensor hand to the diffusion model after the PNG round trip)
__global__ void __launch_bounds__(256)
lidar_pack_kernel(int64_t n_pix, const float* __restrict__ rgb, const float* __restrict__ alpha,
                  uint8_t* __restrict__ out_u8, float* __restrict__ out_f32) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_pix) return;
    float v[4] = {rgb[3 * i], rgb[3 * i + 1], rgb[3 * i + 2], alpha[i]};
    uint32_t q[4];
#pragma unroll
    for (int c = 0; c < 4; ++c) {
        const float t = v[c] * 255.0f;
        q[c] = (uint32_t)(int)fminf(fmaxf(t, 0.0f), 255.0f);  // truncation, as the C cast of astype
    }
    if (out_u8) reinterpret_cast<uint32_t*>(out_u8)[i] = q[0] | (q[1] << 8) | (q[2] << 16) | (q[3] << 24);
    if (out_f32) {
#pragma unroll
        for (int c = 0; c < 4; ++c) out_f32[(int64_t)c * n_pix + i] = (float)q[c] / 255.0f;
    }
}

}  // namespace

extern "C" int64_t sc_lidar_segment_bytes(void) { return (int64_t)sizeof(LidarSeg); }

extern "C" int sc_lidar_gather(int64_t n_points, int n_segs, const void* segs_dev, const float* w2c_row2,
                               float scale, float fx, int height, int width, int use_ndc_scale, float* means,
                               float* rgb, float* scales, void* stream) {
    SC_CHECK_ARG(n_points >= 0 && n_segs >= 0 && n_segs <= LIDAR_MAX_SEGS, "sizes (at most 512 segments)");
    if (n_points == 0) return SC_OK;
    SC_CHECK_ARG(segs_dev && means && rgb && scales && n_segs > 0, "null pointer");
    SC_CHECK_ARG(!use_ndc_scale || w2c_row2, "w2c_row2 needed for the NDC scale");
    const float mn = (float)(height <= width ? height : width);
    lidar_gather_kernel<<<(unsigned)sc_div_up(n_points, 256), 256, 0, (cudaStream_t)stream>>>(
        n_points, n_segs, reinterpret_cast<const LidarSeg*>(segs_dev), w2c_row2, scale, fx, mn, use_ndc_scale, means,
        rgb, scales);
    SC_CHECK_LAUNCH("lidar_gather_kernel");
    return SC_OK;
}

extern "C" int sc_lidar_pack(int64_t n_pix, const float* rgb, const float* alpha, uint8_t* out_u8, float* out_f32,
                             void* stream) {
    SC_CHECK_ARG(n_pix >= 0, "n_pix");
    if (n_pix == 0) return SC_OK;
    SC_CHECK_ARG(rgb && alpha && (out_u8 || out_f32), "null pointer");
    SC_CHECK_ARG(!out_u8 || ((uintptr_t)out_u8 & 3) == 0, "4-byte alignment of out_u8");
    lidar_pack_kernel<<<(unsigned)sc_div_up(n_pix, 256), 256, 0, (cudaStream_t)stream>>>(n_pix, rgb, alpha, out_u8,
                                                                                       out_f32);
    SC_CHECK_LAUNCH("lidar_pack_kernel");
    return SC_OK;
}
