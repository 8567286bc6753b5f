/*
 * har_b200_debug.h — measurement hooks exported by libhar_b200.so next to the reference-facing surface of
 * har_b200.h.  bench.py, the profile scripts and two tests use them; no entry point here has a counterpart in the
 * reference (Main/main.py only takes wall-clock times around fit / transform, main.py:116-118, 299-301, 480-482).
 */
#ifndef HAR_B200_DEBUG_H
#define HAR_B200_DEBUG_H

#include "har_b200.h"

#ifdef __cplusplus
extern "C" {
#endif

/* Kernels launched by this process since the last reset (bench.py's "gpu_launches"). */
long long har_debug_launch_count(int reset);

/* Exchanges issued by har_forest_fit since the last reset: calls, bytes, and whether a library-owned NCCL
 * communicator exists (1) or every exchange went through the caller's callback (0). */
void har_debug_comm_stats(long long* calls, long long* bytes, int* via_nccl, int reset);

/* CUDA-event timing of selected kernels, taken inside the library on the launching stream (off by default).
 * id: 0 node_hist, 1 best_split, 2 part_count, 3 part_scatter, 4 window_features, 5 seg_tiles, 6 predict,
 * 7 bin_features, 8 find_splits, 9 bag_weights.  read() returns accumulated milliseconds, launches and work units. */
void har_debug_profile_enable(int on);
void har_debug_profile_reset(void);
int har_debug_profile_read(int id, double* ms, long long* launches, double* units);

/* Shared-memory atomic micro-benchmark (the yardstick beside k_node_hist): atomics per second, whole device.
 * mode 0: every lane its own bank; mode 1: uniformly random counters in a tile of `words` (power of two <= 8192). */
int har_debug_smem_atomic_peak(int mode, int words, double* atomics_per_s, har_stream_t stream);

/* Mismatches between the count division of k_best_split and __ddiv_rn over n pseudo-random operand pairs. */
int har_debug_div_check(long long n, unsigned long long seed, unsigned long long* mismatches, har_stream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* HAR_B200_DEBUG_H */
