// render.cu — CUDA ray caster for the procedural forest ("needs synthetic image rendering (procedural
// textures rasterised IN CUDA, no graphics API) to have inputs").  The reference reads 960x600 camera images from a rosbag
// (stereo_slam.py:35, :185-205) that is not available offline; the stereo-correspondence stage (Frame::ComputeStereoMatches,
// ORB_SLAM3/src/Frame.cc:811-981; cv2.StereoSGBM, stereo_slam.py:108-123) needs image pairs, so this kernel renders them:
// one thread per pixel casts a ray against the ground plane and the trunk cylinders of forest_slam_b200/synth.py's world
// and shades the hit with a solid (3-D, integer-hashed) texture.  Output: u8 grey image and, optionally, the pixel's true
// disparity fx * baseline / Z_cam as float32 (a dense [H, W] map for the lookup of stereo_slam.py:274-281).
// Only + - * / sqrt floor in fp64, no FMA contraction (-fmad=false): the image is a pure function of the inputs and is
// bit-identical to the NumPy restatement in oracle/render_numpy.py (tests/test_orb_gpu.py).
#include "common.cuh"
#include "../../include/forestslam_b200.h"

namespace {

__device__ __forceinline__ uint32_t hash3(int x, int y, int z) {
    uint32_t h = (uint32_t)x * 0x8da6b343u ^ (uint32_t)y * 0xd8163841u ^ (uint32_t)z * 0xcb1ab31fu;
    h ^= h >> 13;
    h *= 0x5bd1e995u;
    h ^= h >> 15;
    return h;
}

// solid texture at a world point: two octaves of cell noise (5 cm and 40 cm cells), 0..255
__device__ __forceinline__ int solid_texture(double X, double Y, double Z) {
    const int fx = (int)floor(X * 20.0), fy = (int)floor(Y * 20.0), fz = (int)floor(Z * 20.0);
    const int cx = (int)floor(X * 2.5), cy = (int)floor(Y * 2.5), cz = (int)floor(Z * 2.5);
    const int a = (int)(hash3(fx, fy, fz) & 255u), b = (int)(hash3(cx, cy, cz) & 255u);
    return (a * 3 + b * 5) >> 3;
}

struct RenderParams {
    double R[9];   // camera-to-world rotation (row-major), from Frame.T_wc
    double o[3];   // camera centre in the world
    double fx, fy, cx, cy;
    double ground_y;      // world y of the ground plane (y points down; synth.make_world cam_height)
    double fx_baseline;   // fx * baseline (disparity = fx_baseline / Z_cam)
    int width, height, n_trunks;
};

__global__ void __launch_bounds__(256)
k_render_forest(RenderParams P, const double* __restrict__ trunks /* [n][4]: x, z, radius, height */,
                uint8_t* __restrict__ image, float* __restrict__ disparity) {
    const int ix = blockIdx.x * 16 + (threadIdx.x & 15), iy = blockIdx.y * 16 + (threadIdx.x >> 4);
    if (ix >= P.width || iy >= P.height) return;
    const double dcx = ((double)ix - P.cx) / P.fx, dcy = ((double)iy - P.cy) / P.fy;
    const double dx = P.R[0] * dcx + P.R[1] * dcy + P.R[2];
    const double dy = P.R[3] * dcx + P.R[4] * dcy + P.R[5];
    const double dz = P.R[6] * dcx + P.R[7] * dcy + P.R[8];
    double tbest = 1e30;
    int kind = 0;  // 0 sky, 1 ground, 2 trunk
    if (dy > 1e-12) {
        const double t = (P.ground_y - P.o[1]) / dy;
        if (t > 0.0 && t < tbest) {
            tbest = t;
            kind = 1;
        }
    }
    const double a = dx * dx + dz * dz;
    for (int i = 0; i < P.n_trunks; i++) {
        const double ox = P.o[0] - trunks[4 * i], oz = P.o[2] - trunks[4 * i + 1], r = trunks[4 * i + 2];
        const double b = 2.0 * (ox * dx + oz * dz), c = ox * ox + oz * oz - r * r;
        const double disc = b * b - 4.0 * a * c;
        if (disc < 0.0 || a < 1e-18) continue;
        const double t = (-b - sqrt(disc)) / (2.0 * a);
        if (t <= 0.0 || t >= tbest) continue;
        const double y = P.o[1] + t * dy;
        if (y > P.ground_y || y < P.ground_y - trunks[4 * i + 3]) continue;
        tbest = t;
        kind = 2;
    }
    int v;
    float disp = 0.0f;
    if (kind == 0) {
        v = 200 + (iy * 40) / P.height;  // sky: a smooth vertical ramp (no texture: the SAD search rejects it)
    } else {
        const double X = P.o[0] + tbest * dx, Y = P.o[1] + tbest * dy, Z = P.o[2] + tbest * dz;
        const int tex = solid_texture(X, Y, Z);
        v = kind == 1 ? 30 + (tex * 5) / 8 : 60 + (tex * 3) / 4;
        // Z_cam = t (the camera-frame direction has z = 1)
        disp = (float)(P.fx_baseline / tbest);
    }
    image[(size_t)iy * P.width + ix] = (uint8_t)(v > 255 ? 255 : v);
    if (disparity) disparity[(size_t)iy * P.width + ix] = disp;
}

}  // namespace

extern "C" int fs_render_forest(const double* T_wc_host, const FSCamera* cam, double baseline, double ground_y,
                                const double* trunks, int n_trunks, int width, int height, uint8_t* image,
                                float* disparity, fs_stream_t stream) {
    FS_REQUIRE(width > 0 && height > 0 && n_trunks >= 0, "bad size");
    FS_REQUIRE(T_wc_host && cam && image && (trunks || n_trunks == 0), "null pointer");
    RenderParams P;
    for (int i = 0; i < 3; i++) {
        for (int j = 0; j < 3; j++) P.R[3 * i + j] = T_wc_host[4 * i + j];
        P.o[i] = T_wc_host[4 * i + 3];
    }
    P.fx = cam->fx;
    P.fy = cam->fy;
    P.cx = cam->cx;
    P.cy = cam->cy;
    P.ground_y = ground_y;
    P.fx_baseline = cam->fx * baseline;
    P.width = width;
    P.height = height;
    P.n_trunks = n_trunks;
    dim3 grid((width + 15) / 16, (height + 15) / 16);
    k_render_forest<<<grid, 256, 0, (cudaStream_t)stream>>>(P, trunks, image, disparity);
    FS_LAUNCH_CHECK("k_render_forest");
    return FS_OK;
}
