// Internal declarations shared by the host code and the kernels (not part of the ABI).
#pragma once
#include <cstdint>
#include <string>
#include <vector>

#include "../../include/meterelf_b200.h"

// Static per-dial data, built once per params on the host and uploaded as-is.
// The per-dial working window is 64 rows x 64 columns of the dials rect, row r held as
// a 64-bit word (bit c = column win_x0 + c); it covers the dial's mask bounding box plus
// the 2-pixel halo the 3x3 closing needs.
struct MelfDialDev {
    int32_t win_x0, win_y0;      // window origin in dials-rect coordinates (may be < 0)
    int32_t win_w, win_h;        // used extent of the window (<= 64)
    int32_t patch_x, patch_y;    // top-left of the 5x5 colour patch (_reading.py:155-157)
    int32_t range_h, range_l, range_s;
    int32_t neg_mom;
    int32_t n_ring, ring_off;    // slice of the ring tables below
    double aoz_turns;            // angle_of_zero / 360.0 (_reading.py:95)
    uint64_t colvalid;           // bit c: column win_x0+c lies inside the dials rect
    uint64_t rowvalid;           // bit r: row win_y0+r lies inside the dials rect
    uint64_t disc[64];           // DialData.mask (frame pixels of the rect included: findContours pads)
    uint64_t ring[64];           // DialData.circle_mask
    double termx[64];            // sgn(dx)*dx^2 for dx = (win_x0+c) - center_x (_reading.py:36)
    double termy[64];            // same for rows (_reading.py:37)
    // Pixels whose inRange bit can reach the closed mask inside disc | ring: the 5x5 dilation of
    // disc | ring, clipped to the dials rect.  Row r contributes the columns row_c0[r] ..
    // row_c0[r] + (row_base[r+1] - row_base[r]) - 1; flat index = row_base[r] + (c - row_c0[r]).
    int32_t n_need, need_off;    // slice of the need table (one (row << 8) | col entry per flat index)
    uint16_t row_base[65];
    uint8_t row_c0[64];
};

// Ring tables: the pixels of circle_mask sorted by (angle, squared distance), which is
// the order sorted() produces at _reading.py:89 up to the "- 1" unwrap; see DESIGN.md.
struct MelfRingTables {
    std::vector<uint16_t> pos;   // (row << 8) | col inside the window; also raster order key
    std::vector<double> angle;   // get_angle_by_vector(pixel - center)  (_utils.py:18-42)
    std::vector<double> sqd;     // x**2 + y**2                          (_reading.py:66)
    std::vector<uint16_t> need;  // (row << 8) | col of every needed window pixel, raster order, all dials
};

// Host-only (no CUDA): returns MELF_OK or an error code with a message in err.
int melf_host_build_dial(const melf_params *P, int d, MelfDialDev *out,
                         MelfRingTables *rings, std::string *err);
int melf_host_check_params(const melf_params *P, std::string *err);
double melf_host_angle_by_vector(double x, double y);
