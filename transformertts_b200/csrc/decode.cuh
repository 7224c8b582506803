// Argument block of the persistent AR decode kernel (decode.cu); filled by c_api.cu.
#pragma once
#include "kernels.h"

namespace tts {

struct DecLayerW {
  const bf16* w_qkv; const float* b_qkv;   // [768][256]  self-attention in-proj, q rows pre-scaled by 1/8
  const bf16* w_so;  const float* b_so;    // K-blocked transpose [32][256][8] of the self-attention out-proj
  const float* ln1_g; const float* ln1_b;
  const bf16* w_cq;  const float* b_cq;    // [256][256]  cross-attention q proj (pre-scaled)
  const bf16* w_co;  const float* b_co;    // K-blocked transpose [32][256][8] of the cross-attention out-proj
  const float* ln2_g; const float* ln2_b;
  const bf16* w_f1;  const float* b_f1;    // [1024][256]
  const bf16* w_f2;  const float* b_f2;    // K-blocked transpose [128][256][8] of the FFN second matrix
  const float* ln3_g; const float* ln3_b;
};

struct DecodeArgs {
  DecLayerW layer[3];
  const bf16* w_pre1; const float* b_pre1; // [256][128]  (80 mel bins zero-padded to 128)
  const bf16* w_pre2; const float* b_pre2; // K-blocked transpose [32][256][8] of the prenet second matrix
  const bf16* w_head; const float* b_head; // [81][256]   rows 0..79 mel head, row 80 stop head
  const float* pe; float alpha;            // [5000][256]
  const bf16* cross_kv;                    // [3][B][Tp][512]  K cols 0..255, V cols 256..511
  bf16* self_kv;                           // [3][B][Tmax][512]
  const int* ph_len;                       // [B]
  int B, Tp, Tmax;                         // Tmax = max frames (= max_len - 1)
  float stop_logit;                        // logit(stop_threshold)
  float* mel_out;                          // [B][Tmax][80] fp32
  float* stop_out;                         // [B][Tmax]
  int* stop_count;                         // [Tmax], zeroed before launch
  int* n_frames;                           // [1]
  const float* forced_mel;                 // optional [B][Tmax][80]: frame t-1 of this trajectory is fed back at step t
                                           // instead of the kernel's own prediction (teacher-forced decode)
};

int decode_max_clusters(int rows_per_cluster, int cluster_size);   // co-resident clusters (occupancy query)
void decode_plan(int B, int* rows_per_cluster, int* cluster_size);  // one-wave plan with the lowest measured cost
int decode_capacity();                                              // utterances one launch holds co-resident
const char* decode_launch(const DecodeArgs& a, int rows_per_cluster, int cluster_size, cudaStream_t s);

}  // namespace tts
