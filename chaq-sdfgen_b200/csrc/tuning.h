// tuning.h -- the library's tuning switches. They come from the environment, read ONCE when the library is loaded
// (not in the launch path); chaqsdf_reload_tuning() re-reads them (a hook for tests and A/B runs, include/chaqsdf.h).
#pragma once

namespace chaq {

struct Tuning {
    long long band_rows = 0;   // CHAQSDF_BAND_ROWS   rows per band of the general pass 2 (0: automatic)
    long long hull_pb = 0;     // CHAQSDF_HULL_PB     blocks per build run (0: automatic)
    long long hull_sub = 0;    // CHAQSDF_HULL_SUB    1 | 4 sub-runs per build CTA (0: automatic)
    long long hull_eb = 0;     // CHAQSDF_HULL_EB     blocks per paint run (0: automatic)
    long long hull_parts = 0;  // CHAQSDF_HULL_PARTS  column parts of pass 2 that run on separate streams (0: automatic, 1: none)
    bool hull_seq = false;     // CHAQSDF_HULL_SEQ    sequential junction repair (one launch) instead of rounds
    bool merge_plain = false;  // CHAQSDF_MERGE_PLAIN junction rounds without the shared-memory window of hull words
    bool row8_wide = false;    // CHAQSDF_ROW8_WIDE   never use the several-rows-per-warp pass 1 of narrow images
    long long pipe_rows = 0;   // CHAQSDF_PIPE_ROWS   band height of the host pipeline of chaqsdf_u8 (0: automatic)
    long long pipe_first = 0;  // CHAQSDF_PIPE_FIRST  height of its first and last band (0: automatic)
    int dt2d_mode = 0;         // CHAQSDF_DT2D        0 automatic, 1 "stock" (float32 recurrence), 2 "exact"
    int paint = 0;             // CHAQSDF_PAINT       0 automatic (table kernel when it applies), 1 "level": the level-tracking kernel
    int glyph = 0;             // CHAQSDF_GLYPH       0 automatic, 1 never use the fused small-image kernel, 2 always (when it fits)
};

const Tuning& tuning();

}  // namespace chaq
