// Interface between the C-ABI entry point (qgemv_stream.cu: b200_qgemv_tiled) and the stream-K GEMV (qgemv_v4.cu).
#pragma once
#include "common.cuh"

namespace b200 {

// returns B200_OK, an error (message set), or B200_ERR_UNSUPPORTED *without* a message when the shape is outside
// what this kernel covers (the caller then uses the older row-tile kernel)
int qgemv_v4_launch(const b200_qgemv_tiled_args* a, int debug, cudaStream_t st, bool dry = false);  // dry: plan check only
size_t qgemv_v4_workspace_bytes();
// development knobs.  split: -1 auto, 0 whole entities per CTA, 1 stream-K split; xreg: 0 = activations always from shared memory
void qgemv_v4_tune(int split, int xreg);
// development: per-CTA phase time stamps (globaltimer) of the next `slots` launches into buf[slots][160][8] (u64); NULL = off
void qgemv_v4_set_trace(void* buf, int slots);
// development: the trace slot of the next launch (NULL when tracing is off or the slots are used up)
unsigned long long* qgemv_v4_take_trace_slot();
void qgemv_v4_tune_occ2(int on);
void qgemv_v4_tune_kvhint(int on);

// Batched kernel (qgemv_nb.cu): 2..16 plain activation rows, one row tile per warp.  Same return convention as qgemv_v4_launch;
// ws = this kernel's own workspace region (qgemv_nb_workspace_bytes(), zero-initialised once, self-cleaning).
int qgemv_nb_launch(const b200_qgemv_tiled_args* a, uint8_t* ws, unsigned long long* trace, cudaStream_t st, bool dry, int debug = 0);
size_t qgemv_nb_workspace_bytes();
void qgemv_nb_tune_ru(int ru_max);

}  // namespace b200
