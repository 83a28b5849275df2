// svtf_internal.h — shared between the host logic (svtf_host.cpp) and the CUDA side (svtf_api.cu).
#pragma once
#include <cstdint>
#include "../../include/svtfilt_b200.h"

namespace svtf {

// Deblock per-frame tables, passed to the kernel by value (960 bytes).
struct LfTables {
    uint8_t lvl[3][8][2][8][2]; // [plane][segment][dir][ref_frame0][mode_lf]
    uint8_t lim[64], mblim[64], hev[64];
};

void build_lf_tables(const svtf_lf_params &lf, LfTables &t);

// Records the message for svtf_last_error(NULL) and returns `code`.
int set_global_error(int code, const char *msg);

} // namespace svtf
