// warp_kernels.cuh — types shared by the forward-warp kernels (included by tcwarp.cu after the device helpers).
#pragma once

enum { MODE_F32 = 0 };      // accumulation mode tag of the float-reduction kernel (warp_pair.cuh)

// Per-unit header in the workspace, as float bit patterns (uint order == float order for values >= 0).
struct UnitStats {
    unsigned lbits;    // soft-z normaliser: max log(1 + clamp(z', 0, 1000)) over the unit (or the call)
    unsigned zabs;     // max |z'| (written by the float-reduction mode's pre-pass)
    int pad[4];
};
