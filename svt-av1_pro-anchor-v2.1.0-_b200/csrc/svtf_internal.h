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

// ---- loop-restoration filter search, host half (svtf_lrs_host.cpp) ----
// finer_tile_search_wiener_seg (Encoder/Codec/EbRestorationPick.c:1042) as a resumable state machine: while pending(), filter
// the unit with vf / hf, measure the squared error and feed() it; afterwards vf / hf / err hold the refined result.
struct LrsRefine {
    int16_t vf[8], hf[8];
    int64_t err = 0;
    bool    started = false;
    void    start(int win, bool use_refinement, bool one_step, const int16_t *v, const int16_t *h);
    void    feed(int64_t e);
    bool    pending() const { return waiting; }

  private:
    int  win = 7, phase = 2, step = 4, dir = 0, tap = 0, sign = -1;
    bool refine = false, one_step = false, moved_down = false, waiting = false;
    bool propose();
    void next_slot(bool went_down);
    void advance();
};
// finer_search_pixel_proj_error (:320) as a state machine over the two projection coefficients of one (unit, ep): while
// pending(), measure get_pixel_proj_error with xqd and feed() it.
struct LrsSgRefine {
    int     xqd[2] = {0, 0};
    int64_t err    = 0;
    void    start(int ep, bool do_refine, const int init_xqd[2]);
    void    feed(int64_t e);
    bool    pending() const { return waiting; }
    bool    started_err() const { return phase != 0; } // err holds a measured value (the first trial has been fed)

  private:
    int  r[2] = {0, 0}, phase = 2, step = 2, p = 0, sign = -1;
    bool refine = false, moved_down = false, waiting = false;
    bool propose();
    void next_slot(bool went_down);
    void advance();
};
// svt_get_proj_subspace_c (:417) + encode_xq (:508) from the exact integer sums {f0 f0, f1 f1, f0 f1, f0 s, f1 s} of a unit
void lrs_sg_initial_xqd(int ep, const long long sums[5], int size, int xqd[2]);
int  lrs_sgr_bits(int ep, const int xqd[2], const int ref_xqd[2]); // count_sgrproj_bits :655
int    lrs_wiener_bits(int win, const int16_t *vf, const int16_t *hf, const int16_t *ref_vf, const int16_t *ref_hf); // count_wiener_bits :1008
double lrs_rdcost(int rdmult, int64_t rate, int64_t dist);                                                          // RDCOST_DBL

// Records the message for svtf_last_error(NULL) and returns `code`.
int set_global_error(int code, const char *msg);

} // namespace svtf
