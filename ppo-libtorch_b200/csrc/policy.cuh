// policy.cuh — shared-memory tile layout and the actor / critic forward of one tile of TM rows,
// used by the fused policy forward (policy.cu), the fused rollout (rollout.cu) and the fused
// update (update.cu).  Agent::getValue / getActionAndValue* (PPO/Agent.cpp:107-170).
#pragma once
#include "categorical.cuh"
#include "mlp.cuh"

namespace ppo {

// Offsets in floats into dynamic shared memory.
struct TileLayout {
    int a1, a2, ws, xs, ls, vs, end;
    int ldo;  // logits row stride (odd => conflict-free row-per-thread access)
};
__host__ __device__ inline TileLayout tile_layout(int TM, int H, int O, int K) {
    TileLayout L;
    const int lda = lda_of(H);
    L.ldo = K | 1;
    L.a1 = 0;
    L.a2 = L.a1 + TM * lda;
    L.ws = L.a2 + TM * lda;
    L.xs = L.ws + 64 * WS_LD;
    L.ls = L.xs + ((TM * O + 3) & ~3);
    L.vs = L.ls + ((TM * L.ldo + 3) & ~3);
    L.end = L.vs + TM;
    return L;
}

// Forward of one net on the tile: Xs -> a1 -> a2 -> out.  All threads of the CTA must call.
template <int TM, int H>
__device__ inline void net_forward(float* sm, const TileLayout& L, int O, const float* __restrict__ params,
                                   const NetOff& off, int K3, float* out, int ldo) {
    layer1_fwd<TM, H>(sm + L.xs, O, params + off.W1, params + off.b1, sm + L.a1);
    layer2_fwd<TM, H>(sm + L.a1, params + off.W2, params + off.b2, sm + L.a2, sm + L.ws);
    layer3_fwd<TM, H>(sm + L.a2, params + off.W3, params + off.b3, K3, out, ldo);
}

}  // namespace ppo
