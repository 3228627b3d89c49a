// cull.h -- host arithmetic of the whole-pixel culling in ptx_render (api.cu); separate so that the CPU tests can exercise it.
#pragma once
#include <algorithm>
#include <cmath>

#include "../../include/ptx.h"

namespace ptx {

struct CullRect { int x0, y0, x1, y1; };

// Screen rectangle (pixels, half-open) that contains the projection of the box [bmin, bmax] through the pinhole camera, with 2 pixels
// of margin; false when no rectangle can be given (a corner at or behind the camera plane, a degenerate camera).  Double precision:
// a point P is seen at (u, v) with P - O = s ((llc - O) + u hor + v ver), s > 0  (Sensor::get_ray, sensor.h:31-33).
inline bool cull_rect(const ptx_camera& cam, const float* bmin, const float* bmax, int w, int h, CullRect& out) {
    const double O[3] = {cam.origin[0], cam.origin[1], cam.origin[2]};
    double A[3], H[3], V[3];
    for (int k = 0; k < 3; k++) { A[k] = (double)cam.lower_left_corner[k] - O[k]; H[k] = cam.horizontal[k]; V[k] = cam.vertical[k]; }
    auto det3 = [](const double* a, const double* b, const double* c) {
        return a[0] * (b[1] * c[2] - b[2] * c[1]) - b[0] * (a[1] * c[2] - a[2] * c[1]) + c[0] * (a[1] * b[2] - a[2] * b[1]);
    };
    const double D = det3(A, H, V);
    if (!(std::fabs(D) > 1e-30)) return false;
    double umin = 1e300, umax = -1e300, vmin = 1e300, vmax = -1e300;
    for (int corner = 0; corner < 8; corner++) {
        double q[3];
        for (int k = 0; k < 3; k++) {
            const double b = (corner >> k) & 1 ? bmax[k] : bmin[k];
            if (!(b == b) || std::fabs(b) > 1e30) return false;
            q[k] = b - O[k];
        }
        const double s = det3(q, H, V) / D, su = det3(A, q, V) / D, sv = det3(A, H, q) / D;
        if (!(s > 1e-9)) return false;
        const double u = su / s, v = sv / s;
        umin = std::min(umin, u); umax = std::max(umax, u); vmin = std::min(vmin, v); vmax = std::max(vmax, v);
    }
    if (!(umin <= umax) || !(vmin <= vmax)) return false;
    const double x0 = std::floor(umin * w) - 2.0, x1 = std::ceil(umax * w) + 2.0, y0 = std::floor(vmin * h) - 2.0, y1 = std::ceil(vmax * h) + 2.0;
    out.x0 = (int)std::min(std::max(x0, 0.0), (double)w); out.x1 = (int)std::min(std::max(x1, 0.0), (double)w);
    out.y0 = (int)std::min(std::max(y0, 0.0), (double)h); out.y1 = (int)std::min(std::max(y1, 0.0), (double)h);
    if (out.x1 < out.x0) out.x1 = out.x0;
    if (out.y1 < out.y0) out.y1 = out.y0;
    return true;
}

}  // namespace ptx
