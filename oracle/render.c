/*
 * render.c -- oracle (TEST INFRASTRUCTURE ONLY) synthetic frame source: the IARC floor grid
 * of params/map.yaml (10 x 10 lines at 0.32 m pitch spanning +-1.44 m, y = +1.44 green,
 * y = -1.44 red, colours of Mantis3Params.h:40-42) seen by a fisheye (Kannala-Brandt)
 * camera at pose w2c.  It plays the role of the reference's GridRenderer
 * (include/mantis3/GridRenderer.cpp:211-353), whose OpenGL path cannot run here; unlike it,
 * the projection uses the camera intrinsics (the reference left its projection at identity,
 * GridRenderer.cpp:306-317) and the map's 0.32 m pitch (the reference draws 0.3, :232).
 * One sample per pixel centre, ray-cast onto the plane z = 0.
 */
#include "mantis_oracle.h"
#include <math.h>
#include <stddef.h>

void mo_render_grid(const double* R_w2c, const double* t_w2c, const float* K, const double* D, int w, int h,
                    double line_width, int floor_level, uint8_t* bgr)
{
    const double half = 1.44, pitch = 0.32, hw = line_width / 2;
    for (int v = 0; v < h; ++v)
        for (int u = 0; u < w; ++u) {
            uint8_t* o = bgr + ((size_t)v * w + u) * 3;
            o[0] = o[1] = o[2] = (uint8_t)floor_level;
            double px[2] = {(double)u, (double)v}, n[2];
            mo_fisheye_undistort(px, 1, K, D, n);
            if (n[0] == -1000000.0) continue;
            double d[3] = {n[0], n[1], 1.0}, dw[3];
            for (int i = 0; i < 3; ++i) dw[i] = R_w2c[i * 3] * d[0] + R_w2c[i * 3 + 1] * d[1] + R_w2c[i * 3 + 2] * d[2];
            if (!(dw[2] < 0) || !(t_w2c[2] > 0)) continue; /* looking away from the floor */
            double s = -t_w2c[2] / dw[2];
            double X = t_w2c[0] + s * dw[0], Y = t_w2c[1] + s * dw[1];
            if (fabs(X) > half + hw || fabs(Y) > half + hw) continue;
            /* nearest grid line in each direction */
            double kx = floor((X + half) / pitch + 0.5), ky = floor((Y + half) / pitch + 0.5);
            double lx = -half + kx * pitch, ly = -half + ky * pitch;
            int on_x = fabs(X - lx) <= hw, on_y = fabs(Y - ly) <= hw;
            if (!on_x && !on_y) continue;
            if (on_y && ky == 9) { o[0] = 50; o[1] = 255; o[2] = 85; }       /* green edge, y = +1.44 */
            else if (on_y && ky == 0) { o[0] = 50; o[1] = 85; o[2] = 255; }  /* red edge, y = -1.44 */
            else { o[0] = o[1] = o[2] = 255; }
        }
}
