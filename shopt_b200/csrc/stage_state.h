// stage_state.h -- the few integers the stage's device-side control flow carries between kernels (select.cu,
// api_stage.cu): survivor counts, the training split, the current training-list length (ping-pong with the
// compacted lists), the state of the running radix select.  Lives in the stage arena; read back once into pinned memory.
#pragma once
#include "select_core.h"

namespace shopt {

struct StageState {
    long long n_black, n_failed, n_surv;   // whole catalogue (after the all-reduce of the survivor counters)
    long long n_train, cutoff;             // size of the training split; permutation position after its last star
    long long nc;                          // training stars alive, whole catalogue (device copy of the host's count)
    long long nl[2];                       // training stars alive in THIS shard: nl[b] belongs to list buffer b
    long long n_rej;                       // entries dropped by the last compaction (this shard)
    long long n_nan;                       // pixels whose V'P is NaN (NaN -> median repair, interpolate.jl:126-130)
    double thr;                            // sigma clipping: mean + 3 std
    SelState sel;
};

}  // namespace shopt
