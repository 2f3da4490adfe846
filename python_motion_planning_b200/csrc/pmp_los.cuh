// pmp_los.cuh -- the reference's two line-of-sight predicates on the occupancy bitmask, shared by
// pmp_los.cu (batched pairs, path shortcutting) and pmp_bestfirst.cu (Theta* / Lazy Theta* in
// reference order).
#pragma once

#include "pmp_common.cuh"

// theta_star.py:110-171.  The reference evaluates this on Python floats with tau = (d_y - d_x)/2;
// all quantities are integers or half-integers, so e and tau are kept doubled.
__device__ static bool los2d(const GridDev &G, int x1, int y1, int x2, int y2) {
    if (x1 < 0 || y1 < 0 || x2 < 0 || y2 < 0 || x1 >= G.X || x2 >= G.X || y1 >= G.Y || y2 >= G.Y) return false;
    if (cell_blocked(G, x1, y1, 0) || cell_blocked(G, x2, y2, 0)) return false;
    const int d_x = abs(x2 - x1), d_y = abs(y2 - y1);
    const int s_x = (x2 > x1) - (x2 < x1), s_y = (y2 > y1) - (y2 < y1);
    int x = x1, y = y1, e2 = 0;
    int guard = 4 * (d_x + d_y) + 8;
    if (d_x > d_y) {
        const int tau2 = d_y - d_x;
        while (x != x2) {
            if (e2 > tau2) { x += s_x; e2 -= 2 * d_y; }
            else if (e2 < tau2) { y += s_y; e2 += 2 * d_x; }
            else { x += s_x; y += s_y; e2 += 2 * (d_x - d_y); }
            if (cell_blocked(G, x, y, 0) || --guard < 0) return false;
        }
    } else {
        const int tau2 = d_x - d_y;
        while (y != y2) {
            if (e2 > tau2) { y += s_y; e2 -= 2 * d_x; }
            else if (e2 < tau2) { x += s_x; e2 += 2 * d_y; }
            else { x += s_x; y += s_y; e2 += 2 * (d_y - d_x); }
            if (cell_blocked(G, x, y, 0) || --guard < 0) return false;
        }
    }
    return true;
}

// theta_star3d.py:139-215: integer Bresenham along the dominant axis (ties: x, then y)
__device__ static bool los3d(const GridDev &G, int x0, int y0, int z0, int x1, int y1, int z1) {
    if (x0 < 0 || y0 < 0 || z0 < 0 || x0 >= G.X || y0 >= G.Y || z0 >= G.Z) return false;
    if (x1 < 0 || y1 < 0 || z1 < 0 || x1 >= G.X || y1 >= G.Y || z1 >= G.Z) return false;
    if (cell_blocked(G, x0, y0, z0) || cell_blocked(G, x1, y1, z1)) return false;
    const int dx = abs(x1 - x0), dy = abs(y1 - y0), dz = abs(z1 - z0);
    const int sx = x1 >= x0 ? 1 : -1, sy = y1 >= y0 ? 1 : -1, sz = z1 >= z0 ? 1 : -1;
    int x = x0, y = y0, z = z0;
    if (dx >= dy && dx >= dz) {
        int ey = dx / 2, ez = dx / 2;
        while (x != x1) {
            x += sx; ey -= dy; ez -= dz;
            if (ey < 0) { y += sy; ey += dx; }
            if (ez < 0) { z += sz; ez += dx; }
            if (cell_blocked(G, x, y, z)) return false;
        }
        return true;
    }
    if (dy >= dx && dy >= dz) {
        int ex = dy / 2, ez = dy / 2;
        while (y != y1) {
            y += sy; ex -= dx; ez -= dz;
            if (ex < 0) { x += sx; ex += dy; }
            if (ez < 0) { z += sz; ez += dy; }
            if (cell_blocked(G, x, y, z)) return false;
        }
        return true;
    }
    int ex = dz / 2, ey = dz / 2;
    while (z != z1) {
        z += sz; ex -= dx; ey -= dy;
        if (ex < 0) { x += sx; ex += dz; }
        if (ey < 0) { y += sy; ey += dz; }
        if (cell_blocked(G, x, y, z)) return false;
    }
    return true;
}

__device__ static inline bool los(const GridDev &G, int ax, int ay, int az, int bx, int by, int bz) {
    return G.ndim == 2 ? los2d(G, ax, ay, bx, by) : los3d(G, ax, ay, az, bx, by, bz);
}
