// Interface between tomography.cu (C ABI, dispatch) and tomo_group.cu (the group kernel).
#pragma once
#include "common.cuh"

namespace sq {

struct TomoGroupArgs {
    const double* V;      // [B, d]
    double* est;          // [B, d]
    double* n_used;       // [B]
    double* err;          // [B]
    double* ws;           // ctas * G * tomo_group_wsv(L) doubles
    long long B;
    int d, L, logL;
    int G;                // vectors per group
    int norm_inf;
    int attempt;          // index of the attempt in the schedule (keys the Philox streams 2*attempt, 2*attempt + 1)
    double N;             // measurements of the attempt
    unsigned long long seed, vec_offset;
};

// the long-vector variant: members are (vector, leaf block) pairs of tomography.cu's block path
struct TomoBlockGroupArgs {
    const double* V;
    double* est;
    const double* sumsq;   // [B]
    const double* bcnt;    // [B, nblk] measurements that fall in the block (from the top kernel)
    double* bm_a;          // [B, nblk] out, stage 1: '+' mass of the block
    double* bm_b;          // [B, nblk] out, stage 1: '-' mass
    double* lbp;           // [B, nblk] out, stage 1: sum / max of (|x| - P)^2
    double* errp;          // [B, nblk] out, stage 2: error partial
    const int* done;       // [B]
    const double* nhop;    // [B]
    const int* skip2;      // [B]
    double* gws;           // ctas * G * 2048 doubles
    long long B;
    int d, nblk, blk_lo, nown;   // blocks [blk_lo, blk_lo + nown) of every vector
    int G, stage, at, norm_inf, may_skip;
    double N;
    unsigned long long seed, vec_offset;
};
int tomo_blockgroup_launch(TomoBlockGroupArgs a, int ctas, cudaStream_t stream);

// doubles of workspace per vector: the sum tree / counts [2L] and the partial sums of the ordered reductions [L/2]
__host__ __device__ inline size_t tomo_group_wsv(int L) { return (size_t)2 * L + (L / 2 < 2 ? 2 : L / 2); }
size_t tomo_group_ws_bytes(int L, int G, int ctas);
int tomo_group_size(long long B, int L, int sms, int threads);
int tomo_group_launch(TomoGroupArgs a, int ctas, int threads, cudaStream_t stream);

}  // namespace sq
