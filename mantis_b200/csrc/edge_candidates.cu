// edge_candidates.cu -- K1: fused BGR->gray->blur->Sobel->L1 magnitude->NMS->double threshold.
//
// Realises, for a batch of frames, the arithmetic of
//   cv::cvtColor(BGR2GRAY)           include/mantis3/QuadDetection.h:209
//   cv::GaussianBlur(3x3, sigma=3)   :211
//   cv::Canny(T, 3T, 3) up to (not including) hysteresis   :212
// and emits two bit planes per frame (32 px per word, row-major):
//   weak   : pixel is a Canny candidate (mag > low and passes non-max suppression)
//   strong : candidate and mag > high
// The hysteresis closure and the morphology run on-chip in edge_link.cu.
//
// Integer spec (pinned against cv2 4.13 by the oracle, oracle/image_stage.c):
//   gray = (3735 B + 19235 G + 9798 R + 16384) >> 15
//   blur = taps (84,88,84)/256 per axis, BORDER_REFLECT_101, (v + 32768) >> 16
//   Sobel 3x3 on the blurred image with BORDER_REPLICATE; mag = |dx| + |dy|, 0 outside
//   NMS: TG22 = 13573 (Q15) sector test, see nms_pass().
#include "ctx.h"

namespace {

constexpr int TW = 128;  // output tile width  (4 words)
constexpr int TH = 32;   // output tile height
constexpr int GW = TW + 8;  // gray / blurred tile: x0-4 .. x0+TW+4
constexpr int GH = TH + 6;  // y0-3 .. y0+TH+3
constexpr int MW = TW + 2;  // magnitude tile: x0-1 .. x0+TW+1
constexpr int MH = TH + 2;
constexpr int NT = 256;

__device__ __forceinline__ int reflect101(int i, int n)
{
    if (i < 0) i = -i;
    if (i >= n) i = 2 * n - 2 - i;
    return min(max(i, 0), n - 1);  // far outside (never used by in-image outputs)
}

__global__ void __launch_bounds__(NT)
k_edge_candidates(const uint8_t* __restrict__ bgr, int W, int H, int stride, size_t frame_pitch,
                  int low, int high, uint32_t* __restrict__ weak, uint32_t* __restrict__ strong, int WW)
{
    __shared__ uint8_t s_gray[GH][GW];
    __shared__ uint8_t s_blur[GH][GW];
    __shared__ int16_t s_dx[MH][MW];
    __shared__ int16_t s_dy[MH][MW];
    __shared__ uint16_t s_mag[MH][MW];

    const int x0 = blockIdx.x * TW, y0 = blockIdx.y * TH;
    const uint8_t* img = bgr + (size_t)blockIdx.z * frame_pitch;
    const size_t plane = (size_t)blockIdx.z * H * WW;
    const int tid = threadIdx.x;

    // pass 1: gray tile (reflect-101 addressing outside the image)
    for (int i = tid; i < GH * GW; i += NT) {
        int ly = i / GW, lx = i - ly * GW;
        int sy = reflect101(y0 - 3 + ly, H), sx = reflect101(x0 - 4 + lx, W);
        const uint8_t* p = img + (size_t)sy * stride + 3 * sx;
        s_gray[ly][lx] = (uint8_t)((3735 * p[0] + 19235 * p[1] + 9798 * p[2] + 16384) >> 15);
    }
    __syncthreads();

    // pass 2: blurred tile over x0-3..x0+TW+3, y0-2..y0+TH+2.  Positions outside the image
    // take the value at the clamped coordinate (what BORDER_REPLICATE feeds Sobel).
    for (int i = tid; i < (GH - 2) * (GW - 2); i += NT) {
        int ly = i / (GW - 2) + 1, lx = i - (ly - 1) * (GW - 2) + 1;
        int cy = min(max(y0 - 3 + ly, 0), H - 1) - (y0 - 3);
        int cx = min(max(x0 - 4 + lx, 0), W - 1) - (x0 - 4);
        // the clamped coordinate can fall outside this tile only for positions that
        // no in-image output depends on; keep the index legal
        cy = min(max(cy, 1), GH - 2);
        cx = min(max(cx, 1), GW - 2);
        uint32_t h0 = 84u * (s_gray[cy - 1][cx - 1] + s_gray[cy - 1][cx + 1]) + 88u * s_gray[cy - 1][cx];
        uint32_t h1 = 84u * (s_gray[cy][cx - 1] + s_gray[cy][cx + 1]) + 88u * s_gray[cy][cx];
        uint32_t h2 = 84u * (s_gray[cy + 1][cx - 1] + s_gray[cy + 1][cx + 1]) + 88u * s_gray[cy + 1][cx];
        s_blur[ly][lx] = (uint8_t)((84u * (h0 + h2) + 88u * h1 + 32768u) >> 16);
    }
    __syncthreads();

    // pass 3: Sobel + L1 magnitude over x0-1..x0+TW+1, y0-1..y0+TH+1 (0 outside the image)
    for (int i = tid; i < MH * MW; i += NT) {
        int my = i / MW, mx = i - my * MW;
        int gy = y0 - 1 + my, gx = x0 - 1 + mx;
        int dx = 0, dy = 0, m = 0;
        if (gx >= 0 && gx < W && gy >= 0 && gy < H) {
            int ly = my + 2, lx = mx + 3;  // same position in the s_blur frame
            int a = s_blur[ly - 1][lx - 1], b = s_blur[ly - 1][lx], c = s_blur[ly - 1][lx + 1];
            int d = s_blur[ly][lx - 1], f = s_blur[ly][lx + 1];
            int g = s_blur[ly + 1][lx - 1], hh = s_blur[ly + 1][lx], k = s_blur[ly + 1][lx + 1];
            dx = (c - a) + 2 * (f - d) + (k - g);
            dy = (g + 2 * hh + k) - (a + 2 * b + c);
            m = abs(dx) + abs(dy);
        }
        s_dx[my][mx] = (int16_t)dx;
        s_dy[my][mx] = (int16_t)dy;
        s_mag[my][mx] = (uint16_t)m;
    }
    __syncthreads();

    // pass 4: non-max suppression + double threshold; one warp per 32-px row segment
    const int lane = tid & 31, warp = tid >> 5;
    for (int seg = warp; seg < TH * (TW / 32); seg += NT / 32) {
        int ty = seg / (TW / 32), tw = seg - ty * (TW / 32);
        int gy = y0 + ty, gx = x0 + tw * 32 + lane;
        int my = ty + 1, mx = tw * 32 + lane + 1;
        bool cand = false, strg = false;
        if (gx < W && gy < H) {
            int m = s_mag[my][mx];
            if (m > low) {
                int xs = s_dx[my][mx], ys = s_dy[my][mx];
                int ax = abs(xs), ay = abs(ys) << 15;
                int tg22x = ax * 13573;
                if (ay < tg22x) {
                    cand = (m > s_mag[my][mx - 1]) && (m >= s_mag[my][mx + 1]);
                } else {
                    int tg67x = tg22x + (ax << 16);
                    if (ay > tg67x) {
                        cand = (m > s_mag[my - 1][mx]) && (m >= s_mag[my + 1][mx]);
                    } else {
                        int s = ((xs ^ ys) < 0) ? -1 : 1;
                        cand = (m > s_mag[my - 1][mx - s]) && (m > s_mag[my + 1][mx + s]);
                    }
                }
                strg = cand && (m > high);
            }
        }
        uint32_t wb = __ballot_sync(0xffffffffu, cand);
        uint32_t sb = __ballot_sync(0xffffffffu, strg);
        int wj = (x0 >> 5) + tw;
        if (lane == 0 && gy < H && wj < WW) {
            weak[plane + (size_t)gy * WW + wj] = wb;
            strong[plane + (size_t)gy * WW + wj] = sb;
        }
    }
}

}  // namespace

int launch_edge_candidates(mantis_ctx* c, const uint8_t* d_bgr, int n, int w, int h, int stride,
                           size_t frame_pitch, int low, int high, uint32_t* d_weak, uint32_t* d_strong)
{
    dim3 grid((w + TW - 1) / TW, (h + TH - 1) / TH, n);
    k_edge_candidates<<<grid, NT, 0, c->stream>>>(d_bgr, w, h, stride, frame_pitch, low, high, d_weak,
                                                  d_strong, words_per_row(w));
    c->launches++;
    MCHECK(c, cudaGetLastError());
    return MANTIS_OK;
}
