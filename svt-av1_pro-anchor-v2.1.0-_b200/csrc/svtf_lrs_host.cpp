// svtf_lrs_host.cpp — host half of the loop-restoration filter search (Wiener and self-guided tools).  Pure C++ (no CUDA), part of
// libsvtfilt_b200.so; the per-sample work (statistics, trial filtering + SSE) is in svtf_lrs.cuh / svtf_lr.cuh.
//
// Reference (relative to Source/Lib/): Encoder/Codec/EbRestorationPick.c — wiener_decompose_sep_sym :906 with
// update_a_sep_sym :805 / update_b_sep_sym :857 / linsolve_wiener :766, finalize_sym_filter :977, compute_score :937,
// finer_tile_search_wiener_seg :1042 (here a resumable state machine, so that the trials of all units of a picture run in
// lock-step, one launch per round), count_wiener_bits :1008 over svt_aom_count_primitive_refsubexpfin
// (Encoder/Codec/EbEntropyCoding.c:2876-3022), search_wiener_finish :1380 and the frame decision of rest_finish_search :1555.
// All arithmetic is 64-bit integer with C's truncating division, in the reference's order of operations; the two RD
// comparisons are IEEE double (RDCOST_DBL, Common/Codec/EbRestoration.h:346).
#include <algorithm>
#include <climits>
#include <cmath>
#include <cstdint>
#include <cstdlib>
#include <cstring>

#include "svtf_internal.h"

namespace svtf {

namespace {
constexpr int     kWin = 7, kHalf = 3, kStep = 128 /* WIENER_FILT_STEP */, kIters = 5 /* NUM_WIENER_ITERS */;
constexpr int64_t kScale = int64_t(1) << 16; /* WIENER_TAP_SCALE_FACTOR */
// WIENER_FILT_TAPn_{MINV, MAXV, SUBEXP_K} and the initial (mid) taps incl. the centre (Common/Codec/EbRestoration.h:131-153)
const int kTapMin[3] = {-5, -23, -17}, kTapMax[3] = {10, 8, 46}, kTapK[3] = {1, 2, 3}, kInit[7] = {3, -7, 15, 106, 15, -7, 3};

inline int fold(int i, int win) { return i > (win >> 1) ? win - 1 - i : i; } // wrap_index :759

// linsolve_wiener: n <= 3 unknowns, row stride `ld`
bool solve(int n, int64_t *A, int ld, int64_t *b, int32_t *x) {
    for (int k = 0; k + 1 < n; k++) {
        for (int i = n - 1; i > k; i--) {
            int64_t *lo = A + (i - 1) * ld, *hi = A + i * ld;
            if (std::llabs(lo[k]) < std::llabs(hi[k])) {
                std::swap_ranges(lo, lo + n, hi);
                std::swap(b[i - 1], b[i]);
            }
        }
        const int64_t *piv = A + k * ld;
        for (int i = k + 1; i < n; i++) {
            if (piv[k] == 0)
                return false;
            int64_t      *row = A + i * ld;
            const int64_t c = row[k], cd = piv[k];
            for (int j = 0; j < n; j++) row[j] -= c / 256 * piv[j] / cd * 256;
            b[i] -= c * b[k] / cd;
        }
    }
    for (int i = n - 1; i >= 0; i--) {
        const int64_t *row = A + i * ld;
        if (row[i] == 0)
            return false;
        int64_t c = 0;
        for (int j = i + 1; j < n; j++) c += row[j] * x[j] / kScale;
        x[i] = (int32_t)(kScale * (b[i] - c) / row[i]);
    }
    return true;
}

// One half-step of the alternating least squares: the filter of one direction is held (`held`), the other solved.
// solve_rows = false: update_a_sep_sym (b held, vertical taps solved); true: update_b_sep_sym.
void half_step(int win, const int64_t *M, const int64_t *H, const int32_t *held, bool solve_rows, int32_t *out) {
    const int win2 = win * win, n1 = (win >> 1) + 1, c = n1 - 1;
    int64_t   rhs[kHalf + 1] = {}, mat[(kHalf + 1) * (kHalf + 1)] = {};
    for (int i = 0; i < win; i++)
        for (int j = 0; j < win; j++) {
            // M[i * win + j]: i = column offset of the window, j = row offset
            const int     dst = solve_rows ? fold(i, win) : fold(j, win);
            const int32_t w   = solve_rows ? held[j] : held[i];
            rhs[dst] += M[i * win + j] * w / kScale;
        }
    for (int i = 0; i < win; i++)
        for (int j = 0; j < win; j++)
            for (int k = 0; k < win; k++)
                for (int l = 0; l < win; l++) {
                    if (solve_rows)
                        mat[fold(j, win) * n1 + fold(i, win)] += H[(i * win + k) * win2 + j * win + l] * held[k] / kScale * held[l] / kScale;
                    else
                        mat[fold(l, win) * n1 + fold(k, win)] += H[(j * win + k) * win2 + i * win + l] * held[i] / kScale * held[j] / kScale;
                }
    // the taps sum to one: substitute the centre tap
    const int64_t rc = rhs[c], mcc = mat[c * n1 + c];
    for (int i = 0; i < c; i++) rhs[i] -= rc * 2 + mat[i * n1 + c] - 2 * mcc;
    for (int i = 0; i < c; i++)
        for (int j = 0; j < c; j++) mat[i * n1 + j] -= 2 * (mat[i * n1 + c] + mat[c * n1 + j] - 2 * mcc);
    int32_t s[kWin];
    if (!solve(c, mat, n1, rhs, s))
        return; // keep the previous taps
    s[c] = (int32_t)kScale;
    for (int i = n1; i < win; i++) {
        s[i] = s[win - 1 - i];
        s[c] -= 2 * s[i];
    }
    std::memcpy(out, s, sizeof(int32_t) * win);
}

// finalize_sym_filter: scaled taps -> the coded InterpKernel (centre tap stored without its implicit +128)
void to_kernel(int win, const int32_t *f, int16_t *k) {
    std::memset(k, 0, sizeof(int16_t) * 8);
    for (int i = 0; i < (win >> 1); i++) {
        const int64_t num = (int64_t)f[i] * kStep;
        k[i]              = (int16_t)((num < 0 ? num - kScale / 2 : num + kScale / 2) / kScale);
    }
    auto clip = [](int v, int t) { return (int16_t)std::min(std::max(v, kTapMin[t]), kTapMax[t]); };
    if (win == kWin) {
        k[0] = clip(k[0], 0), k[1] = clip(k[1], 1), k[2] = clip(k[2], 2);
    } else { // shorter windows are stored one slot further in, from the first two solved taps
        k[2] = clip(k[1], 2), k[1] = clip(k[0], 1), k[0] = 0;
    }
    k[6] = k[0], k[5] = k[1], k[4] = k[2];
    k[3] = (int16_t)(-2 * (k[0] + k[1] + k[2]));
}

// compute_score: x'Hx - 2 x'M of the separable filter minus the same for the identity filter
int64_t filter_score(int win, const int64_t *M, const int64_t *H, const int16_t *vf, const int16_t *hf) {
    const int win2 = win * win, off = (kWin - win) >> 1;
    int16_t   a[kWin], b[kWin];
    a[kHalf] = b[kHalf] = kStep;
    for (int i = 0; i < kHalf; i++) {
        a[i] = a[kWin - 1 - i] = vf[i], b[i] = b[kWin - 1 - i] = hf[i];
        a[kHalf] = (int16_t)(a[kHalf] - 2 * a[i]), b[kHalf] = (int16_t)(b[kHalf] - 2 * b[i]);
    }
    int32_t ab[kWin * kWin];
    for (int k = 0; k < win; k++)
        for (int l = 0; l < win; l++) ab[k * win + l] = a[l + off] * b[k + off];
    int64_t lin = 0, quad = 0;
    for (int k = 0; k < win2; k++) {
        lin += ab[k] * M[k] / kStep / kStep;
        for (int l = 0; l < win2; l++) quad += ab[k] * H[k * win2 + l] * ab[l] / kStep / kStep / kStep / kStep;
    }
    const int mid = win2 >> 1;
    return (quad - 2 * lin) - (H[mid * win2 + mid] - 2 * M[mid]);
}

int recentre(int r, int v) { return v > 2 * r ? v : (v >= r ? 2 * (v - r) : 2 * (r - v) - 1); }
// svt_aom_count_primitive_refsubexpfin(n, k, ref, v)
int subexp_bits(int n, int k, int ref, int v) {
    v = 2 * ref <= n ? recentre(ref, v) : recentre(n - 1 - ref, n - 1 - v);
    int bits = 0, mk = 0;
    for (int i = 0;; i++) {
        const int b = i ? k + i - 1 : k, a = 1 << b;
        if (n <= mk + 3 * a) { // svt_aom_count_primitive_quniform(n - mk, v - mk)
            const int nn = n - mk;
            if (nn > 1) {
                int l = 1;
                while ((nn - 1) >> l) l++;
                bits += (v - mk) < (1 << l) - nn ? l - 1 : l;
            }
            return bits;
        }
        bits++;
        if (v < mk + a)
            return bits + b;
        mk += a;
    }
}
} // namespace

int lrs_wiener_bits(int win, const int16_t *vf, const int16_t *hf, const int16_t *ref_vf, const int16_t *ref_hf) {
    int bits = 0;
    for (int t = win == kWin ? 0 : 1; t < 3; t++) bits += subexp_bits(kTapMax[t] - kTapMin[t] + 1, kTapK[t], ref_vf[t] - kTapMin[t], vf[t] - kTapMin[t]);
    for (int t = win == kWin ? 0 : 1; t < 3; t++) bits += subexp_bits(kTapMax[t] - kTapMin[t] + 1, kTapK[t], ref_hf[t] - kTapMin[t], hf[t] - kTapMin[t]);
    return bits;
}

double lrs_rdcost(int rdmult, int64_t rate, int64_t dist) { return ((double)rate * rdmult) / (double)(1 << 9) + (double)dist * (1 << 7); }

// ---- finer_tile_search_wiener_seg as a state machine ----------------------------------------------------------------
// Usage: start(); while (pending()) { <filter the unit with vf/hf, measure> ; feed(err); }  -> best error in `err`.
void LrsRefine::start(int win_, bool use_refinement, bool one_step_, const int16_t *v, const int16_t *h) {
    win = win_, refine = use_refinement, one_step = one_step_;
    std::memcpy(vf, v, sizeof(vf)), std::memcpy(hf, h, sizeof(hf));
    started = true, phase = 0, waiting = true; // the first trial: the solved taps as they are
}

// Applies the move of the current (step, filter, tap, sign) if it stays inside the coded range.
bool LrsRefine::propose() {
    int16_t *f = dir ? vf : hf;
    if (sign < 0 ? f[tap] - step < kTapMin[tap] : f[tap] + step > kTapMax[tap])
        return false;
    f[tap] = (int16_t)(f[tap] + sign * step), f[kWin - 1 - tap] = (int16_t)(f[kWin - 1 - tap] + sign * step);
    f[kHalf] = (int16_t)(f[kHalf] - 2 * sign * step);
    return true;
}

// Moves on to the next (sign, tap, filter, step) after the current sign is finished; `went_down`: the downward sweep of
// this tap lowered the error, which ends the tap loop of this filter (the `if (skip) break;` of the reference).
void LrsRefine::next_slot(bool went_down) {
    if (sign < 0 && !went_down) {
        sign = 1;
        return;
    }
    sign = -1;
    if (!went_down && tap + 1 < kHalf) {
        tap++;
        return;
    }
    tap = (kWin - win) >> 1;
    if (dir == 0) {
        dir = 1;
        return;
    }
    dir = 0;
    step >>= 1;
    if (step < (one_step ? 4 : 1))
        phase = 2;
}

void LrsRefine::advance() { // find the next trial or finish
    while (phase == 1) {
        if (propose()) {
            waiting = true;
            return;
        }
        next_slot(moved_down);
        moved_down = false;
    }
    waiting = false;
}

void LrsRefine::feed(int64_t e) {
    waiting = false;
    if (phase == 0) {
        err = e;
        if (!refine) {
            phase = 2;
            return;
        }
        phase = 1, step = 4, dir = 0, tap = (kWin - win) >> 1, sign = -1, moved_down = false;
        advance();
        return;
    }
    int16_t *f = dir ? vf : hf;
    if (e > err) { // undo the move, this sign is finished
        f[tap] = (int16_t)(f[tap] - sign * step), f[kWin - 1 - tap] = (int16_t)(f[kWin - 1 - tap] - sign * step);
        f[kHalf] = (int16_t)(f[kHalf] + 2 * sign * step);
        next_slot(moved_down);
        moved_down = false;
    } else {
        err = e;
        if (sign < 0)
            moved_down = true;
        if (!(step == 4 && !one_step)) { // only the largest step keeps moving in the same direction
            next_slot(moved_down);
            moved_down = false;
        }
    }
    advance();
}

// ---- self-guided tools ------------------------------------------------------------------------------------------------
namespace {
const int kSgR[16][2] = {{2, 1}, {2, 1}, {2, 1}, {2, 1}, {2, 1}, {2, 1}, {2, 1}, {2, 1}, {2, 1}, {2, 1}, {0, 1}, {0, 1}, {0, 1}, {0, 1}, {2, 0}, {2, 0}};
const int kSgMin[2] = {-96, -32}, kSgMax[2] = {31, 95}; // SGRPROJ_PRJ_MIN0/1, MAX0/1 (Common/Codec/EbRestoration.h:101-104)
inline int clampi(int v, int lo, int hi) { return v < lo ? lo : (v > hi ? hi : v); }
} // namespace

// The reference accumulates the five sums in double; every term is an integer below 2^34 and the totals stay below 2^53,
// so its sums are exact and equal the integer sums converted once.  Everything after that is the reference's double
// arithmetic, operation by operation.
void lrs_sg_initial_xqd(int ep, const long long sums[5], int size, int xqd[2]) {
    const int r0 = kSgR[ep][0], r1 = kSgR[ep][1];
    double    h00 = (double)sums[0], h11 = (double)sums[1], h01 = (double)sums[2], c0 = (double)sums[3], c1 = (double)sums[4];
    h00 /= size, h01 /= size, h11 /= size, c0 /= size, c1 /= size;
    const double h10 = h01;
    int          xq[2] = {0, 0};
    if (r0 == 0) {
        if (!(h11 < 1e-8))
            xq[1] = (int)std::rint(c1 / h11 * 128);
    } else if (r1 == 0) {
        if (!(h00 < 1e-8))
            xq[0] = (int)std::rint(c0 / h00 * 128);
    } else {
        const double det = h00 * h11 - h01 * h10;
        if (!(det < 1e-8)) {
            const double x0 = (h11 * c0 - h01 * c1) / det, x1 = (h00 * c1 - h10 * c0) / det;
            xq[0] = (int)std::rint(x0 * 128), xq[1] = (int)std::rint(x1 * 128);
        }
    }
    if (r0 == 0)
        xqd[0] = 0, xqd[1] = clampi(128 - xq[1], kSgMin[1], kSgMax[1]);
    else if (r1 == 0)
        xqd[0] = clampi(xq[0], kSgMin[0], kSgMax[0]), xqd[1] = clampi(128 - xqd[0], kSgMin[1], kSgMax[1]);
    else
        xqd[0] = clampi(xq[0], kSgMin[0], kSgMax[0]), xqd[1] = clampi(128 - xqd[0] - xq[1], kSgMin[1], kSgMax[1]);
}

int lrs_sgr_bits(int ep, const int xqd[2], const int ref_xqd[2]) {
    int bits = 4; // SGRPROJ_PARAMS_BITS
    for (int p = 0; p < 2; p++)
        if (kSgR[ep][p] > 0)
            bits += subexp_bits(kSgMax[p] - kSgMin[p] + 1, 4 /* SGRPROJ_PRJ_SUBEXP_K */, ref_xqd[p] - kSgMin[p], xqd[p] - kSgMin[p]);
    return bits;
}

void LrsSgRefine::start(int ep, bool do_refine, const int init_xqd[2]) {
    r[0] = kSgR[ep][0], r[1] = kSgR[ep][1];
    refine = do_refine, xqd[0] = init_xqd[0], xqd[1] = init_xqd[1];
    phase = 0, waiting = true;
}
bool LrsSgRefine::propose() {
    if (sign < 0 ? xqd[p] - step < kSgMin[p] : xqd[p] + step > kSgMax[p])
        return false;
    xqd[p] += sign * step;
    return true;
}
void LrsSgRefine::next_slot(bool went_down) {
    if (sign < 0 && !went_down) {
        sign = 1;
        return;
    }
    sign = -1;
    if (!went_down) // next coefficient that has a filter behind it
        for (int q = p + 1; q < 2; q++)
            if (r[q] > 0) {
                p = q;
                return;
            }
    step >>= 1;
    if (step < 1) {
        phase = 2;
        return;
    }
    p = r[0] > 0 ? 0 : 1;
}
void LrsSgRefine::advance() {
    while (phase == 1) {
        if (propose()) {
            waiting = true;
            return;
        }
        next_slot(moved_down);
        moved_down = false;
    }
    waiting = false;
}
void LrsSgRefine::feed(int64_t e) {
    waiting = false;
    if (phase == 0) {
        err = e;
        if (!refine) {
            phase = 2;
            return;
        }
        phase = 1, step = 2, p = r[0] > 0 ? 0 : 1, sign = -1, moved_down = false;
        advance();
        return;
    }
    if (e > err) {
        xqd[p] -= sign * step;
        next_slot(moved_down);
        moved_down = false;
    } else {
        err = e;
        if (sign < 0)
            moved_down = true;
        if (step != 2) { // only the first step size keeps moving in the same direction
            next_slot(moved_down);
            moved_down = false;
        }
    }
    advance();
}

} // namespace svtf

// search_wiener_seg :1296 from the statistics on: 1 = taps usable, 0 = the reference keeps RESTORE_NONE for the unit
extern "C" int svtf_lr_wiener_from_stats(int32_t win, const int64_t *M, const int64_t *H, int16_t vfilter[8], int16_t hfilter[8]) {
    using namespace svtf;
    if ((win != 7 && win != 5 && win != 3) || !M || !H || !vfilter || !hfilter)
        return set_global_error(SVTF_E_ARG, "svtf_lr_wiener_from_stats: bad argument");
    const int off = (kWin - win) >> 1;
    int32_t   a[kWin], b[kWin];
    for (int i = 0; i < win; i++) a[i] = b[i] = (int32_t)(kScale / kStep * kInit[i + off]);
    for (int it = 1; it < kIters; it++) {
        half_step(win, M, H, b, false, a);
        half_step(win, M, H, a, true, b);
    }
    to_kernel(win, a, vfilter), to_kernel(win, b, hfilter);
    return filter_score(win, M, H, vfilter, hfilter) > 0 ? 0 : 1;
}
