/* include/glowtts_b200_debug.h -- instrumentation entry points of libglowtts_b200.so.
 *
 * NOT part of the product boundary: these symbols exist only in a library built with -DGTK_INSTRUMENT
 * (GTK_EXTRA_CFLAGS=-DGTK_INSTRUMENT python -m glowtts_b200.build --force), which also turns on the per-CTA
 * globaltimer stamps inside the tensor-core and MAS kernels.  The default build exports none of them
 * (tests/test_abi.py checks that).  Used by the scripts under profiles/. */
#ifndef GLOWTTS_B200_DEBUG_H
#define GLOWTTS_B200_DEBUG_H
#include <stddef.h>
#ifdef __cplusplus
extern "C" {
#endif

/* Debug: phase timestamps (ns) of the first n_ctas (<= 512) CTAs of the last fused WN-layer launch:
 * out[cta * 6 + {start, setup done, GEMM1 done, epilogue-1 done, GEMM2 done, end}], cta = batch * tiles + tile.
 * Synchronises the device. */
int gtk_debug_wn_trace(unsigned long long* out, int n_ctas);
/* Debug: SM cycles per CTA of the last fused WN-layer launch: out[cta * 4 + {MMA thread waiting for weight stages during
 * GEMM 1, MMA thread GEMM-1 total, TMA producer waiting for free stages, producer total}]. */
int gtk_debug_wn_wait(long long* out, int n_ctas);
/* The same for GEMM b (the k = 5 data gradient) of the fused layer-backward kernel. */
int gtk_debug_wb_wait(long long* out, int n_ctas);
/* Debug: DP-warp phase stamps of the last TMA-path MAS launch: out[cta * 6 + {start, first tile landed, DP done,
 * backtrack done, barrier passed, end}] (globaltimer ns).  Synchronises the device. */
int gtk_debug_mas_trace(unsigned long long* out, int n_ctas);
/* Debug: SM cycles the DP warp of each CTA spent waiting for value tiles after the first one (same launch). */
int gtk_debug_mas_wait(long long* out, int n_ctas);
/* Debug: the fused layer-backward kernel: out[cta * 7 + {start, setup, GEMM a, epilogue a, GEMM b, epilogue b, end}]. */
int gtk_debug_wb_trace(unsigned long long* out, int n_ctas);
/* Debug: the whole-WaveNet kernel (wn_stack.cu): out[cta * 24 + {0 start, 1 prologue done, 2 start-conv epilogue done,
 * per layer l: 3+4l GEMM 1 complete, 4+4l epilogue 1 done, 5+4l GEMM 2 complete, 6+4l epilogue 2 done, 3+4L end}];
 * wait: out[cta * 8 + l] = SM cycles the MMA warp waited for weight stages in GEMM 1 of layer l, [4 + l] = that GEMM 1. */
int gtk_debug_ws_trace(unsigned long long* out, int n_ctas);
int gtk_debug_ws_wait(long long* out, int n_ctas);
/* Debug: stream-K weight-gradient kernel (wgrad_group.cu), SM cycles per CTA of the last launch: out[cta * 8 + {0 MMA warp
 * waiting for operand stages, 1 waiting for the previous segment's drain, 2 MMA warp total, 3 k-steps, 4 segments, 5 kernel
 * entry -> MMA warp start, 6 producer waiting for free stages, 7 drain time of warp 4}]. */
int gtk_debug_gg_wait(long long* out, int n_ctas);
/* Debug: the same for the plain tcgen05 conv kernel.  Select which launches record (epilogue 0 plain / 1 gate-bwd /
 * 2 affine, and the tap count; -1 = none), run, then read out[cta * 5 + {start, setup done, accumulator complete,
 * epilogue math done, stores issued}] of the last selected launch.  Synchronises the device. */
int gtk_debug_conv_trace_select(int epilogue, int taps);
int gtk_debug_conv_trace(unsigned long long* out, int n_ctas);
/* out[cta * 2 + {0: SM cycles the TMA producer waited for a free stage, 1: cycles the MMA thread waited for data}] */
int gtk_debug_conv_wait(long long* out, int n_ctas);
/* Hardware probe (debug_shift.cu): D_s = A[s : s+128] B^T for row shifts s = 1..4 of a K-major SWIZZLE_128B operand,
 * descriptor base-offset field 0 (mode 0) or (addr >> 7) & 7 (mode 1).  a [136][64] bf16, b [64][64] bf16,
 * out [(s-1)*2 + mode][128][64] fp32. */
int gtk_debug_shift_probe(const void* a, const void* b, float* out, void* stream);

#ifdef __cplusplus
}
#endif
#endif
