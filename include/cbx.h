/* chatterbox_b200 C-ABI — the drop-in boundary underneath the reference's Python API.
 *
 * The reference (rossturner/chatterbox) has no FFI layer: its boundary is the Python class API of
 * src/chatterbox/tts.py and src/chatterbox/vc.py.  This library sits under the host-side mirrors
 * of the inner seams the hot path goes through; each entry point cites the reference interface it
 * replaces.  Conventions: plain C structs, raw DEVICE pointers (tensor.data_ptr()), every weight /
 * input / output buffer is caller-owned (PyTorch), int return code (0 = ok) + cbx_last_error(),
 * nothing throws across the ABI, calls are stream-ordered on the cudaStream_t passed as `void*`.
 * There is no CPU fallback: every call fails loudly without a CUDA device.
 */
#ifndef CBX_H
#define CBX_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

const char* cbx_last_error(void);
int cbx_version(void);

/* ------------------------------------------------------------------------------------------
 * (1) T3 0.5B Llama backbone: prefill + per-token decode loop.
 * Replaces T3.inference (src/chatterbox/models/t3/t3.py:203-371) below prepare_input_embeds:
 * T3HuggingfaceBackend.forward (t3/inference/t3_hf_backend.py:73-116) = HF LlamaModel x30 +
 * speech_head, the CFG pair as one batch-2 pass (t3.py:289-295,323-326), temperature /
 * repetition-penalty / min-p / top-p processing (t3.py:331-337), sampling (t3.py:340-341), EOS
 * test and next-token embedding (t3.py:343-356).
 * ------------------------------------------------------------------------------------------ */
typedef struct cbx_t3 cbx_t3;

typedef struct CbxT3Config {
  int32_t device;     /* CUDA ordinal */
  int32_t n_layers;   /* 30 (llama_configs.py:8); smaller only in tests */
  int32_t max_pos;    /* KV slab capacity in positions (cond 34 + text + 2 BOS + speech) */
  int32_t kv_fp32;    /* 0: bf16 KV slab (perf mode), 1: fp32 KV slab (parity mode) */
  int32_t speech_pos_rows; /* rows of speech_pos_emb (4100); decode calls that would index past it are rejected; 0 = unchecked */
} CbxT3Config;

typedef struct CbxT3Weights {          /* device pointers, caller-owned, must outlive the handle */
  const void* wstream;                 /* bf16 per-CTA weight streams packed by the plan below: [G][stride_units][16384] */
  int64_t stride_units;                /* units (32 KB) between consecutive CTA streams */
  const float* ln1;                    /* [n_layers][1024] input_layernorm.weight */
  const float* ln2;                    /* [n_layers][1024] post_attention_layernorm.weight */
  const float* final_norm;             /* [1024] tfmr.norm.weight */
  const float* speech_emb;             /* [8194][1024] speech_emb.weight (t3.py:52) */
  const float* speech_pos_emb;         /* [4100][1024] speech_pos_emb.emb.weight (t3.py:60) */
  const float* rope_cos;               /* [max_pos][32] cos of llama3-scaled RoPE angles, fp32 */
  const float* rope_sin;               /* [max_pos][32] */
  const void* wplain;                  /* optional (may be NULL): the same weights as plain bf16 row-major matrices for the
                                          GEMM prefill — per layer [q|k|v 3072x1024][o 1024x1024][gate/up 8192x1024 with 128
                                          gate rows then the 128 matching up rows per 256-row tile][down 1024x4096], then
                                          speech_head [8194][1024].  NULL: the prompt is prefilled by the decode kernel,
                                          one position per weight pass */
} CbxT3Weights;

typedef struct CbxSampleParams {       /* the generate() knobs of tts.py:208-218 */
  float cfg_weight;                    /* > 0: rows (cond, uncond), logits = c + w (c - u) */
  float temperature;
  float repetition_penalty;
  float min_p;
  float top_p;
  int32_t greedy;                      /* 1: argmax of the processed logits (parity mode); 0: multinomial */
  int32_t stop_on_eos;                 /* stop at token 6562 (t3.py:347) */
  uint64_t seed;                       /* Philox seed for multinomial sampling */
} CbxSampleParams;

/* Number of CTAs (= SMs) the megakernel runs with on `device`; the weight stream is packed for it. */
int cbx_t3_grid(int device);
/* The static work plan (csrc/t3_plan.h): units of phase `ph` (0 qkv,1 o,2 gate/up,3 down,4 head)
 * consumed by CTA `c` of `G`, and the id of the j-th one.  Used by the host-side weight packer. */
int cbx_t3_plan_count(int ph, int c, int G);
int cbx_t3_plan_unit(int ph, int c, int G, int j);
/* k-slices of a weight unit = warps of a megakernel CTA (8 or 16): the packer's fragment order depends on it. */
int cbx_t3_kwarps(void);

int cbx_t3_create(const CbxT3Config* cfg, const CbxT3Weights* w, cbx_t3** out);
int cbx_t3_destroy(cbx_t3* h);
/* Prefill: embeds [B][P][1024] fp32 (output of prepare_input_embeds + doubled BOS, t3.py:240-295),
 * B in {1,2}.  Resets the KV slab and the generated-id set (BOS 6561 seen), leaves the logits of
 * the last position in the handle (t3.py:307-316). */
int cbx_t3_prefill(cbx_t3* h, const float* embeds, int B, int P, void* stream);
/* Decode loop (t3.py:319-367) as ONE persistent launch: up to max_new_tokens tokens; if
 * forced_tokens != NULL the loop is teacher-forced with those ids.  tokens_out [max_new_tokens]
 * int32, n_out [1] int32 (tokens emitted; -(n + 1) if the device watchdog fired), logits_out (nullable)
 * [max_new_tokens][B][8194] raw logits per step.  Calls that would run past the KV slab or the learned
 * position table are rejected with an error before anything is launched. */
int cbx_t3_decode(cbx_t3* h, const CbxSampleParams* sp, int max_new_tokens, const int32_t* forced_tokens,
                  int32_t* tokens_out, int32_t* n_out, float* logits_out, void* stream);
/* Host-visible state after a synchronised stream: cache length, tokens generated, device error flag,
 * and the number of kernel launches issued through this handle. */
int cbx_t3_state(cbx_t3* h, int32_t* pos, int32_t* n_gen, int32_t* err, int64_t* launches);

/* AlignmentStreamAnalyzer (src/chatterbox/models/t3/inference/alignment_stream_analyzer.py:29-154; BASELINE config 5): call
 * before cbx_t3_prefill.  layer >= 0 (the reference uses 9) turns it on for the following utterances: the head-mean attention
 * of the conditional row at that layer over the text positions [text_begin, text_begin + n_text) is taken from the prompt
 * prefill and from every decode step's attention phase, the step() state machine runs on the device before each token is
 * sampled, and the EOS logit (6562) is forced / suppressed as there (:141-153).  layer < 0 turns it off.  Needs the bf16 KV
 * slab, the GEMM prefill (wplain) and max_pos <= 2176; one decode launch per utterance.  Optional device buffers: inject_rows
 * [rows][n_text] replaces the measured attention (tests of the state machine), rows_out [rows][n_text] receives every row
 * consumed (rows = prompt rows + one per step), decisions_out [max_new][4] = {text position, started | complete << 1 |
 * forced << 2 | suppressed << 3, arg-max column, rows so far}.  (The reference ships the call disabled,
 * t3_hf_backend.py:108-109; eos_idx here is the stop-speech token as upstream Chatterbox passes it.) */
int cbx_t3_set_alignment(cbx_t3* h, int layer, int text_begin, int n_text, const float* inject_rows, float* rows_out,
                         int32_t* decisions_out);

/* Debug aid: CTA 0 writes %globaltimer (ns) after every phase of decode iteration `it` into
 * stamps[] (device, >= 512 entries); NULL disables.  The stamps are compiled in only with
 * -DT3_PROFILE (phases) or -DT3_TIMELINE (every CTA, stamps[] >= 1024 + 1024 * SMs entries): in the
 * default build even the dead tests cost issue slots on the critical path, and this call records nothing. */
int cbx_t3_set_profile(cbx_t3* h, uint64_t* stamps, int it);

/* ------------------------------------------------------------------------------------------
 * (1c) T3 conditioning front-end: the prompt embeddings cbx_t3_prefill consumes, once per utterance.
 * Replaces T3CondEnc.forward (src/chatterbox/models/t3/modules/cond_enc.py:64-97), Perceiver / AttentionBlock2
 * (modules/perceiver.py:113-212), the exaggeration projection (cond_enc.py:85-88; tts.py:204,225-231),
 * T3.prepare_conditioning / prepare_input_embeds (t3.py:71-108) and the doubled BOS of t3.py:284-295.  fp32 throughout,
 * like the reference; every input is a DEVICE pointer (the emotion scalar and the token ids are read on the device: no
 * host synchronisation).
 * ------------------------------------------------------------------------------------------ */
typedef struct cbx_t3fe cbx_t3fe;

typedef struct CbxT3FrontWeights {       /* device pointers, fp32, caller-owned (reference state-dict keys in the comments) */
  const float* spkr_w;                   /* cond_enc.spkr_enc.weight [1024][256] */
  const float* spkr_b;                   /* cond_enc.spkr_enc.bias [1024] */
  const float* query;                    /* cond_enc.perceiver.pre_attention_query [32][1024] */
  const float* norm_g; const float* norm_b;              /* cond_enc.perceiver.attn.norm.{weight,bias} */
  const float* wq; const float* bq;      /* cond_enc.perceiver.attn.to_q.{weight [1024][1024], bias} */
  const float* wk; const float* bk;      /* ...to_k */
  const float* wv; const float* bv;      /* ...to_v */
  const float* wo; const float* bo;      /* ...proj_out */
  const float* emotion_w;                /* cond_enc.emotion_adv_fc.weight [1024][1] */
  const float* text_emb;                 /* text_emb.weight [704][1024] */
  const float* text_pos_emb;             /* text_pos_emb.emb.weight [2050][1024] */
  const float* speech_emb;               /* speech_emb.weight [8194][1024] */
  const float* speech_pos_emb;           /* speech_pos_emb.emb.weight [4100][1024] */
} CbxT3FrontWeights;

int cbx_t3_front_create(int device, int max_cond_tokens, const CbxT3FrontWeights* w, cbx_t3fe** out);
int cbx_t3_front_destroy(cbx_t3fe* h);
/* speaker_emb [256]; cond_tokens int64 [n_cond] (T3Cond.cond_prompt_speech_tokens) or, when non-NULL, cond_prompt_emb
 * [n_cond][1024] (T3Cond.cond_prompt_speech_emb) instead; n_cond == 0: no perceiver rows; emotion_adv [1];
 * text_tokens int64 [B][N] incl. SOT / EOT; zero_uncond_text: row 1's text-token embeddings are zeroed (cfg_weight > 0,
 * t3.py:90-92); n_bos: 1, or 2 for the doubled BOS of the CFG path.  embeds_out [B][P][1024], P = 1 + 32 + 1 + N + n_bos
 * (2 + N + n_bos without conditioning tokens). */
int cbx_t3_front(cbx_t3fe* h, const float* speaker_emb, const int64_t* cond_tokens, const float* cond_prompt_emb, int n_cond,
                 const float* emotion_adv, const int64_t* text_tokens, int B, int N, int zero_uncond_text, int n_bos,
                 float* embeds_out, void* stream);
int cbx_t3_front_launches(cbx_t3fe* h, int64_t* launches);

/* ------------------------------------------------------------------------------------------
 * (1b) Batched T3 decode: up to 64 concurrent utterances ("slots") per handle, one pass of all slots' CFG pairs
 * through every weight matrix per step (BASELINE config 4).  Replaces the reference server's one-request-at-a-time
 * loop (src/server/server.py:189-299: one synchronous tts_model.generate per request, :223,232; global
 * model.conds swap, src/server/voice_manager.py:370-379): the per-request state — KV slab, generated-id set,
 * sampling knobs, Philox stream, EOS flag — lives in per-slot device arrays, so requests with different voices and
 * knobs share the 1.02 GB weight stream of a step.  Same arithmetic per utterance as (1): hi / lo activation
 * planes against bf16 weights, fp32 accumulate, bf16 (or fp32) KV, identical logit processing and sampling.
 * ------------------------------------------------------------------------------------------ */
typedef struct cbx_t3b cbx_t3b;

typedef struct CbxT3BatchConfig {
  int32_t device;
  int32_t n_layers;        /* 30 */
  int32_t n_slots;         /* U: concurrent utterances, 1..64 (2 U rows = one 128-row GEMM tile) */
  int32_t max_pos;         /* KV capacity of ONE slot in positions */
  int32_t kv_fp32;         /* 0: bf16 slab, 1: fp32 slab (parity mode) */
  int32_t max_new_cap;     /* token capacity of a slot (max max_new_tokens of a request) */
  int32_t keep_logits;     /* 1: keep raw per-step logits [slot][max_new_cap][2][8194] (tests; 67 KB per token) */
  int32_t speech_pos_rows; /* rows of speech_pos_emb, 0 = unchecked */
} CbxT3BatchConfig;

/* Needs CbxT3Weights.wplain (the plain matrices); wstream is not used by this engine. */
int cbx_t3_batch_create(const CbxT3BatchConfig* cfg, const CbxT3Weights* w, cbx_t3b** out);
int cbx_t3_batch_destroy(cbx_t3b* h);
/* Admit a request into `slot`: prompt prefill of embeds [2][P][1024] (CFG pair; duplicate the row for cfg_weight 0)
 * into the slot's KV slab, pending logits, per-slot knobs.  forced_tokens: nullable DEVICE array [max_new_tokens]
 * (teacher forcing).  The slot then takes part in every following cbx_t3_decode_batch until EOS / max_new_tokens. */
int cbx_t3_batch_prefill(cbx_t3b* h, int slot, const float* embeds, int P, const CbxSampleParams* sp, int max_new_tokens,
                         const int32_t* forced_tokens, void* stream);
/* The same for n requests with prompts of EQUAL length P into the consecutive slots [slot0, slot0 + n): embeds
 * [n][2][P][1024], sp = HOST array [n], forced_tokens = nullable DEVICE array [n][max_new_tokens].  The n prompts go through
 * the layers together (2 n P rows per GEMM instead of 2 P: the reference prefills one request per generate() call,
 * t3.py:307-316, server.py:223-232); results per slot are those of cbx_t3_batch_prefill up to the GEMMs' summation order. */
int cbx_t3_batch_prefill_many(cbx_t3b* h, int slot0, int n, const float* embeds, int P, const CbxSampleParams* sp,
                              int max_new_tokens, const int32_t* forced_tokens, void* stream);
/* n_steps decode steps (t3.py:319-367, one token per running slot and step), each ONE CUDA-graph launch.  slots:
 * HOST array of the slots to advance, NULL = every running slot.  Finished / free slots ride along as inert rows. */
int cbx_t3_decode_batch(cbx_t3b* h, const int32_t* slots, int n_slots, int n_steps, void* stream);
/* Synchronises `stream` and reports per slot (HOST arrays [n_slots], each nullable): tokens emitted, done flag
 * (1 finished, 0 running, -1 free), cache length. */
int cbx_t3_batch_poll(cbx_t3b* h, int32_t* n_gen, int32_t* done, int32_t* pos, void* stream);
/* Copies the first n tokens of `slot` into a DEVICE buffer (stream-ordered). */
int cbx_t3_batch_tokens(cbx_t3b* h, int slot, int32_t* tokens_out, int n, void* stream);
/* keep_logits engines: copies the raw logits of the slot's first n steps, [n][2][8194] fp32, into a DEVICE buffer. */
int cbx_t3_batch_logits(cbx_t3b* h, int slot, float* logits_out, int n, void* stream);
/* Frees the slot for the next request. */
int cbx_t3_batch_release(cbx_t3b* h, int slot, void* stream);
int cbx_t3_batch_info(cbx_t3b* h, int32_t* kv_splits, int64_t* launches, int64_t* launches_per_step);

/* ------------------------------------------------------------------------------------------
 * (2) S3Gen conditional flow matching: the Euler/CFG solve around the ConditionalDecoder estimator.
 * Replaces CausalConditionalCFM.forward / ConditionalCFM.solve_euler
 * (src/chatterbox/models/s3gen/flow_matching.py:194-218, 72-123) and ConditionalDecoder.forward
 * (s3gen/decoder.py:233-317; call signature forward(x, mask, mu, t, spks, cond), flow_matching.py:125-127).
 * Tensors are the reference's channel-first fp32 layouts; the mask is all-ones (one un-padded utterance).
 * ------------------------------------------------------------------------------------------ */
typedef struct cbx_cfm cbx_cfm;

typedef struct CbxCfmWeights {     /* device pointers, caller-owned, must outlive the handle */
  const void* bf16_blob;           /* GEMM matrices, bf16, [N][taps*K] each (chatterbox_b200/cfm.py: pack order) */
  const float* f32_blob;           /* biases, LayerNorm gains, time-embedding MLPs */
  const int64_t* bf16_off;         /* HOST array [270]: element offsets into bf16_blob */
  const int64_t* f32_off;          /* HOST array [502]: element offsets into f32_blob */
} CbxCfmWeights;

int cbx_cfm_create(int device, int max_len, const CbxCfmWeights* w, cbx_cfm** out);
int cbx_cfm_destroy(cbx_cfm* h);
/* One estimator call on a batch of 2 rows sharing t: x, mu, cond [2][80][L], spks [2][80] -> out [2][80][L]. */
int cbx_cfm_estimator(cbx_cfm* h, const float* x, const float* mu, const float* spks, const float* cond, float t,
                      int L, float* out, void* stream);
/* n_steps Euler steps with classifier-free guidance: z, mu, cond [80][L], spks [80]; t_vals / dt_vals are HOST
 * arrays [n_steps] (time fed to step s and its step size, computed by the caller exactly as flow_matching.py
 * does); mel_out [80][L]. */
int cbx_cfm_solve(cbx_cfm* h, const float* z, const float* mu, const float* spks, const float* cond, int L,
                  int n_steps, const float* t_vals, const float* dt_vals, float cfg_rate, float* mel_out, void* stream);
/* Batched variants (BASELINE config 4: the flow-matching stage of up to max_batch utterances as ONE solve, so the estimator's
 * GEMMs see 2 nb L rows instead of 2 L).  z, mu, cond [nb][80][L], spks [nb][80], mel_out [nb][80][L], every utterance padded to
 * L frames; lens: HOST array [nb] of true lengths (NULL = all L).  The padding is the reference's padded-batch semantics
 * (decoder.py:25-33, utils/mask.py:167-193): the estimator is causal in its convolutions, so frames past lens[u] reach valid
 * frames only through attention, where they are masked as keys; output frames past lens[u] are unspecified (the reference
 * zeroes them) and must be sliced off by the caller. */
int cbx_cfm_create_batched(int device, int max_len, int max_batch, const CbxCfmWeights* w, cbx_cfm** out);
int cbx_cfm_solve_batch(cbx_cfm* h, int nb, const float* z, const float* mu, const float* spks, const float* cond,
                        const int32_t* lens, int L, int n_steps, const float* t_vals, const float* dt_vals, float cfg_rate,
                        float* mel_out, void* stream);
int cbx_cfm_launches(cbx_cfm* h, int64_t* launches);
/* Entries of the handle's two bounded caches: executable solve graphs (LRU, at most 8) and per-length op plans (at most 64). */
int cbx_cfm_cache_sizes(cbx_cfm* h, int32_t* n_graphs, int32_t* n_plans);

/* ------------------------------------------------------------------------------------------
 * (2b) S3Gen flow front-end: speech tokens -> (mu, spks, cond), the inputs of the flow-matching solve, once per utterance.
 * Replaces CausalMaskedDiffWithXvec.inference up to the decoder call (src/chatterbox/models/s3gen/flow.py:236-270):
 * F.normalize + spk_embed_affine_layer, input_embedding, UpsampleConformerEncoder.forward
 * (s3gen/transformer/upsample_encoder.py:237-304; rel-pos attention transformer/attention.py:225-330 with rel_shift as an
 * index transform), encoder_proj, and the prompt-mel conditioning tensor.  One un-padded utterance per call.
 * ------------------------------------------------------------------------------------------ */
typedef struct cbx_flowfe cbx_flowfe;

typedef struct CbxFlowFrontWeights {   /* device pointers, caller-owned, must outlive the handle */
  const void* bf16_blob;               /* Linear / Conv1d matrices as bf16 [N][taps * K] (chatterbox_b200/flow_frontend.py: pack order) */
  const float* f32_blob;               /* embedding table, biases, LayerNorm gains, pos_bias_u / v, speaker projection */
  const int64_t* bf16_off;             /* HOST array [56] */
  const int64_t* f32_off;              /* HOST array [115] */
} CbxFlowFrontWeights;

int cbx_flowfe_create(int device, int max_tokens, const CbxFlowFrontWeights* w, cbx_flowfe** out);
int cbx_flowfe_destroy(cbx_flowfe* h);
/* prompt_token int64 [Tp], token int64 [T] (negative ids clamp to 0, flow.py:254), prompt_feat fp32 [Lp][80] (Lp = 2 Tp in the
 * reference), embedding fp32 [192]; outputs for L = 2 (Tp + T) mel frames: mu_out [80][L], spks_out [80], cond_out [80][L]. */
int cbx_flowfe_run(cbx_flowfe* h, const int64_t* prompt_token, int Tp, const int64_t* token, int T, const float* prompt_feat,
                   int Lp, const float* embedding, float* mu_out, float* spks_out, float* cond_out, void* stream);
int cbx_flowfe_launches(cbx_flowfe* h, int64_t* launches);

/* ------------------------------------------------------------------------------------------
 * (3) HiFT-GAN vocoder: mel -> f0 -> harmonic source -> waveform.
 * Replaces HiFTGenerator.inference / decode (src/chatterbox/models/s3gen/hifigan.py:462-474, 412-444),
 * ConvRNNF0Predictor.forward (s3gen/f0_predictor.py:52-55), SourceModuleHnNSF / SineGen (hifigan.py:200-283)
 * and the trim_fade of S3Token2Wav.inference (s3gen/s3gen.py:233-236, 296).  mel is the reference's
 * channel-first fp32 [80][L]; waveforms are [480 L] fp32 at 24 kHz.
 * ------------------------------------------------------------------------------------------ */
typedef struct cbx_hift cbx_hift;

typedef struct CbxHiftWeights {    /* device pointers, caller-owned, must outlive the handle */
  const void* bf16_blob;           /* convolution matrices, weight-norm folded, bf16 (chatterbox_b200/hift.py pack order) */
  const float* f32_blob;           /* biases, Snake alphas, 18-channel source convolutions, small linears */
  const int64_t* bf16_off;         /* HOST array [82] */
  const int64_t* f32_off;          /* HOST array [164] */
} CbxHiftWeights;

int cbx_hift_create(int device, int max_frames, const CbxHiftWeights* w, cbx_hift** out);
int cbx_hift_destroy(cbx_hift* h);
int cbx_hift_f0(cbx_hift* h, const float* mel, int L, float* f0_out, void* stream);
/* phase [9] / noise [9][480 L] (device) may be NULL: then drawn from Philox(seed) (the reference uses torch's global RNG). */
int cbx_hift_source(cbx_hift* h, const float* f0, int L, const float* phase, const float* noise, uint64_t seed,
                    float* s_out, void* stream);
int cbx_hift_decode(cbx_hift* h, const float* mel, int L, const float* s, int apply_fade, float* wav_out, void* stream);
/* f0 -> source (unless source_in is given) -> decode; source_out (nullable) receives the source signal used. */
int cbx_hift_inference(cbx_hift* h, const float* mel, int L, const float* source_in, uint64_t seed, int apply_fade,
                       float* wav_out, float* source_out, void* stream);
int cbx_hift_launches(cbx_hift* h, int64_t* launches);

/* ------------------------------------------------------------------------------------------
 * (3b) Allocation-free harmonic source.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it).
 * Same arithmetic and kernels as cbx_hift_source (SineGen + SourceModuleHnNSF, hifigan.py:200-283), but handle-free
 * and without any allocation: the double-precision phase-prefix scratch is a file-scope static __device__ array of
 * cbx_hift_source_static_max_frames() = 4096 entries (32 KB per GPU, held until exit); larger L is rejected with an
 * error.  The array is shared by all calls on a device: calls that overlap on the GPU race — serialise them on one
 * stream.  lin_w [9] / lin_b [1]: m_source.l_linear.weight / .bias (device, fp32); f0 [L]; phase [9] / noise [9][480 L]
 * nullable (then drawn from Philox(seed)); s_out [480 L].
 * ------------------------------------------------------------------------------------------ */
int cbx_hift_source_static_max_frames(void);
int cbx_hift_source_static(int device, const float* lin_w, const float* lin_b, const float* f0, int L, const float* phase,
                           const float* noise, uint64_t seed, float* s_out, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* CBX_H */
