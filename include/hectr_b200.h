/* hectr_b200 -- C ABI of the B200 (sm_100a) CKKS evaluation engine.
 *
 * Drop-in boundary for the evaluation half of the GPQHE C API that HECTR's encrypted MPC
 * (src/hempc.c:216-274) drives.  Every entry point names the reference function it stands in
 * for (paths relative to the OChicken/HECTR tree).  The reference computes on libgcrypt MPIs
 * in host memory; here a ciphertext lives in GPU memory as fixed-width integers and a handle
 * may hold a BATCH of independent ciphertexts that every operation processes in one launch.
 *
 * Data formats (all little-endian, host or device as stated):
 *   ciphertext  uint64 words[batch][2][W_l][N]   two's-complement coefficient, word w of
 *               coefficient i of polynomial c_h of instance b at ((b*2+h)*W_l + w)*N + i,
 *               W_l = hb_ctx_words(ctx, l) = ceil((log2 q_l + 1)/64).  Same integers as
 *               he_ct_t.c0/c1.coeffs[i] (GPQHE/src/gpqhe.h:84-91), centred in [-q_l/2, q_l/2).
 *   plaintext   uint64 words[WC][N], WC = hb_ctx_words(ctx, L)        (he_pt_t.m, gpqhe.h:93-97)
 *   eval. key   uint64 p0[dimevk][N], p1[dimevk][N] exactly as he_genswk stores them
 *               (NTT/RNS form, GPQHE/src/he-kem.c:103-110; he_evk_t gpqhe.h:78-82)
 *   RNS limbs   uint64 limbs[nlimbs][N]
 *
 * Restrictions, checked at hb_ctx_create: q = 2^logq and Delta = 2^logDelta (every reference call
 * site uses powers of two: src/ctr.c:514-518, GPQHE/tests/gpqhe.c:1349-1353).
 *
 * Error behaviour follows the reference (GPQHE/src/precomp.c:344-350, he-add.c:35): misuse prints
 * "error: ..." on stderr and abort()s; so does any CUDA failure.  There is no CPU fallback.
 * One context is used by one host thread at a time (the reference has one global context).
 */
#ifndef HECTR_B200_H
#define HECTR_B200_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

typedef struct hb_ctx hb_ctx;

typedef struct hb_info {           /* what hectx_init derives (precomp.c:386-409) */
  unsigned logn, n, logq, logDelta, slots;
  unsigned L, dim, dimevk, dimub;
  unsigned words_top;              /* W_L */
  unsigned nbitsP, nbitsPqL;
  int device, sm_count;
} hb_info;

typedef struct hb_ct {             /* he_ct_t (gpqhe.h:84-91) with a device payload */
  unsigned l;
  double nu, B;
  unsigned batch;
  uint64_t *words;                 /* DEVICE pointer, owned by the library */
} hb_ct;

typedef struct hb_pt {             /* he_pt_t (gpqhe.h:93-97) */
  double nu;
  uint64_t *words;                 /* DEVICE pointer [WC][N] */
} hb_pt;

typedef struct hb_evk {            /* he_evk_t (gpqhe.h:78-82) */
  uint64_t *p;                     /* DEVICE pointer [2][dimevk][N] pairs (key limb, its Shoup quotient) */
} hb_evk;

typedef struct hb_sk {             /* secret key, poly_mpi_t sk (he-kem.c:31-50), held as transform limbs */
  uint64_t *hat;                   /* DEVICE [dimub][N]: ntt(sk mod p_d), canonical */
  uint64_t *scaled[64];            /* DEVICE, built on first use: Shoup pairs of hat * N^-1 (P_D/p_d)^-1 for D limbs */
} hb_sk;

/* ---- context: hectx_init / hectx_exit (precomp.c:431-453,465-) --------------------------------- */
hb_ctx *hb_ctx_create(unsigned logn, unsigned logq, unsigned slots, unsigned logDelta, int device);
void hb_ctx_destroy(hb_ctx *ctx);
void hb_ctx_info(const hb_ctx *ctx, hb_info *out);
unsigned hb_ctx_words(const hb_ctx *ctx, unsigned l);          /* W_l */
unsigned hb_ctx_logq(const hb_ctx *ctx, unsigned l);           /* log2 q_l */
void hb_ctx_primes(const hb_ctx *ctx, uint64_t *out);          /* [dimub]       rns_ctx.p */
void hb_ctx_twiddles(const hb_ctx *ctx, unsigned d, int inverse, uint64_t *out); /* [N] plain zeta^brv(i) */
unsigned hb_dim_keyswitch(const hb_ctx *ctx, unsigned l);      /* he-automorphism.c:52 */
unsigned hb_dim_mul(const hb_ctx *ctx, unsigned l);            /* he-mult.c:99 */
unsigned hb_dim_mulpt(const hb_ctx *ctx, unsigned l, double nu); /* he-mult.c:168 */
void hb_ctx_set_stream(hb_ctx *ctx, void *cuda_stream);        /* default: the context's own stream */
void hb_sync(hb_ctx *ctx);
/* CUDA-event stopwatch on the context's stream (bench.py, roofline timing) */
void hb_timer_start(hb_ctx *ctx);
float hb_timer_stop_ms(hb_ctx *ctx);
/* kernels launched by this context since creation (bench.py "gpu_launches") */
uint64_t hb_launch_count(const hb_ctx *ctx);
/* per-kernel CUDA-event timing on the context's stream: enable, run, then read {"kernel": {"launches","ms"}} as JSON
 * (returns the length the full report needs) */
void hb_prof_enable(hb_ctx *ctx, int on);
void hb_prof_reset(hb_ctx *ctx);
size_t hb_prof_report(hb_ctx *ctx, char *buf, size_t cap);

/* ---- host-only view of the same tables (no GPU needed: CPU tests of the host logic) -------------------- */
typedef struct hb_tables hb_tables;
hb_tables *hb_tables_create(unsigned logn, unsigned logq, unsigned slots, unsigned logDelta);  /* NULL + message if rejected */
void hb_tables_destroy(hb_tables *t);
void hb_tables_info(const hb_tables *t, hb_info *out);
void hb_tables_primes(const hb_tables *t, uint64_t *out);
void hb_tables_twiddles(const hb_tables *t, unsigned d, int inverse, uint64_t *out);
/* CRT constants for D limbs: inv[D] = (P_D/p_d)^-1 mod p_d, phat_low[D][W_L], P_low[W_L]   (precomp.c:266-293) */
void hb_tables_crt(const hb_tables *t, unsigned D, uint64_t *inv, uint64_t *phat_low, uint64_t *P_low);
unsigned hb_tables_dim(const hb_tables *t, int which, unsigned l, double nu);   /* 0 keyswitch, 1 mul, 2 mulpt */
void hb_tables_encode(const hb_tables *t, const double *m, double Delta, uint64_t *words);
void hb_tables_decode(const hb_tables *t, const uint64_t *words, unsigned W, double Delta, double *m);

/* ---- raw memory helpers ------------------------------------------------------------------------- */
void *hb_dev_alloc(hb_ctx *ctx, size_t bytes);
void hb_dev_free(hb_ctx *ctx, void *p);
void *hb_host_alloc(hb_ctx *ctx, size_t bytes);                /* pinned */
void hb_host_free(hb_ctx *ctx, void *p);
void hb_h2d(hb_ctx *ctx, void *dst, const void *src, size_t bytes);   /* async on the ctx stream */
void hb_d2h(hb_ctx *ctx, void *dst, const void *src, size_t bytes);   /* async on the ctx stream */
void hb_dev_fill(hb_ctx *ctx, void *dst, int byte, size_t bytes);

/* ---- containers: he-mem.c:31-97 ---------------------------------------------------------------- */
void hb_ct_alloc(hb_ctx *ctx, hb_ct *ct, unsigned batch);      /* he_alloc_ct */
void hb_ct_free(hb_ctx *ctx, hb_ct *ct);                       /* he_free_ct  */
void hb_ct_copy(hb_ctx *ctx, hb_ct *dst, const hb_ct *src);    /* he_copy_ct  */
void hb_ct_upload(hb_ctx *ctx, hb_ct *ct, const uint64_t *host_words, unsigned l, double nu, double B);
void hb_ct_download(hb_ctx *ctx, const hb_ct *ct, uint64_t *host_words);   /* synchronises */
/* device-to-device exchange with buffers the caller owns (e.g. a torch tensor used for the NCCL result gather) */
void hb_ct_export(hb_ctx *ctx, const hb_ct *ct, void *dst_device);
void hb_ct_import(hb_ctx *ctx, hb_ct *ct, const void *src_device, unsigned l, double nu, double B);
void hb_pt_alloc(hb_ctx *ctx, hb_pt *pt);                      /* he_alloc_pt */
void hb_pt_free(hb_ctx *ctx, hb_pt *pt);
void hb_pt_upload(hb_ctx *ctx, hb_pt *pt, const uint64_t *host_words, double nu);
void hb_evk_upload(hb_ctx *ctx, hb_evk *evk, const uint64_t *p0, const uint64_t *p1);  /* he_alloc_evk + copy */
void hb_evk_free(hb_ctx *ctx, hb_evk *evk);

/* ---- encode / decode on the host: he-encode.c:53-125, canemb.c:43-81 -------------------------------- */
/* m: slots complex doubles (re,im interleaved).  words: host [WC][N]; only indices j*gap and j*gap+N/2 are
 * written, exactly like encode() (hazard: other coefficients are left as they are). */
void hb_encode(const hb_ctx *ctx, const double *m, double Delta, uint64_t *words);
void hb_decode(const hb_ctx *ctx, const uint64_t *words, unsigned W, double Delta, double *m);
void hb_he_ecd(hb_ctx *ctx, hb_pt *pt, const double *m);       /* he_ecd: encode at Delta and upload */
void hb_he_const_pt(hb_ctx *ctx, hb_pt *pt, double re, double im);   /* he_const_pt he-encode.c:119-125 */

/* ---- evaluation ------------------------------------------------------------------------------------ */
void hb_he_add(hb_ctx *ctx, hb_ct *ct, const hb_ct *a, const hb_ct *b);        /* he_add   he-add.c:32-54 */
void hb_he_sub(hb_ctx *ctx, hb_ct *ct, const hb_ct *a, const hb_ct *b);        /* he_sub   he-add.c:57-79 */
void hb_he_neg(hb_ctx *ctx, hb_ct *ct);                                        /* he_neg   he-add.c:124-142 */
void hb_he_addpt(hb_ctx *ctx, hb_ct *dst, const hb_ct *src, const hb_pt *pt);  /* he_addpt he-add.c:80-100 */
void hb_he_subpt(hb_ctx *ctx, hb_ct *dst, const hb_ct *src, const hb_pt *pt);  /* he_subpt he-add.c:103-123 */
void hb_he_rs(hb_ctx *ctx, hb_ct *ct);                                         /* he_rs    he-rescale.c:33-54 */
void hb_he_moddown(hb_ctx *ctx, hb_ct *ct);                                    /* he_moddown he-rescale.c:56-70 */
void hb_he_mulpt(hb_ctx *ctx, hb_ct *dst, const hb_ct *src, const hb_pt *pt);  /* he_mulpt he-mult.c:159-196 */
void hb_he_mul(hb_ctx *ctx, hb_ct *ct, const hb_ct *a, const hb_ct *b, const hb_evk *rlk); /* he_mul he-mult.c:88-156 */
void hb_he_rot(hb_ctx *ctx, hb_ct *ct, int rot, const hb_evk *rk);             /* he_rot   he-automorphism.c:102-115; rk[rot] */
void hb_he_conj(hb_ctx *ctx, hb_ct *ct, const hb_evk *ck);                     /* he_conj  he-automorphism.c:87-100 */
/* he_gemv he-algo.c:47-93.  A: slots x slots complex doubles, row-major, shared by the whole batch. */
void hb_he_gemv(hb_ctx *ctx, hb_ct *dst, const double *A, const hb_ct *ct, const hb_evk *rk);
/* The ciphertext half of ctr_hempc (src/hempc.c:253-266):
 *   up = uhat|moddown - ( Ka (xhat - xr) + Kb (uhat - ur) )      all at level l, result at l-1 */
void hb_ctr_hempc_eval(hb_ctx *ctx, hb_ct *up, const double *Ka, const double *Kb, const hb_ct *xhat,
                       const hb_ct *xr, const hb_ct *uhat, const hb_ct *ur, const hb_evk *rk);
/* Same, end to end from HOST buffers (pinned or pageable): uploads the four input batches, evaluates,
 * downloads the result batch (level l-1) and synchronises.  in_*: [batch][2][W_l][N]. */
void hb_ctr_hempc_eval_host(hb_ctx *ctx, unsigned batch, unsigned l, const double *Ka, const double *Kb,
                            const uint64_t *in_xhat, const uint64_t *in_xr, const uint64_t *in_uhat,
                            const uint64_t *in_ur, const hb_evk *rk, uint64_t *out_up);

/* Streaming form for a server that evaluates batch after batch: enqueue upload, evaluation and download of one batch and
 * return a ticket; hb_ctr_hempc_wait blocks until that batch's out_up is complete.  Two submissions may be in flight (two
 * persistent device buffer sets), so the PCIe transfers of batch k+1 and k-1 overlap the evaluation of batch k while every
 * launch keeps the full batch size; a third submission first waits for the first.  Host buffers should be pinned
 * (hb_host_alloc) and must stay valid until the wait returns. */
uint64_t hb_ctr_hempc_submit_host(hb_ctx *ctx, unsigned batch, unsigned l, const double *Ka, const double *Kb,
                                  const uint64_t *in_xhat, const uint64_t *in_xr, const uint64_t *in_uhat,
                                  const uint64_t *in_ur, const hb_evk *rk, uint64_t *out_up);
void hb_ctr_hempc_wait(hb_ctx *ctx, uint64_t ticket);

/* ---- compositions of he-algo.c, built only from the entry points above: hb_algo.cpp -------------- */
void hb_he_sum(hb_ctx *ctx, hb_ct *dst, const hb_ct *ct, const hb_evk *rk);                       /* he_sum  he-algo.c:95-103 */
void hb_he_idx(hb_ctx *ctx, hb_ct *dst, const hb_ct *ct, unsigned idx, const hb_evk *rk);         /* he_idx  :105-112 */
void hb_he_nrm2(hb_ctx *ctx, hb_ct *dst, const hb_ct *ct, const hb_evk *rlk, const hb_evk *ck, const hb_evk *rk); /* he_nrm2 :114-125 */
void hb_he_inv(hb_ctx *ctx, hb_ct *dst, const hb_ct *ct, const hb_evk *rlk, unsigned iter);       /* he_inv  :131-164 */
void hb_he_sqrt(hb_ctx *ctx, hb_ct *dst, const hb_ct *ct, const hb_evk *rlk, unsigned iter);      /* he_sqrt :166-206 */
void hb_he_sigmoid(hb_ctx *ctx, hb_ct *dst, const hb_ct *ct, const hb_evk *rlk);                  /* he_sigmoid :208-277 */
void hb_he_log(hb_ctx *ctx, hb_ct *dst, const hb_ct *ct, const hb_evk *rlk);                      /* he_log  :279-362 */
void hb_he_exp(hb_ctx *ctx, hb_ct *dst, double a_re, double a_im, const hb_ct *ct, const hb_evk *rlk, unsigned iter); /* he_exp :364-458 */
void hb_he_cmp(hb_ctx *ctx, hb_ct *dst, const hb_ct *ct1, const hb_ct *ct2, const hb_evk *rlk, unsigned iter, unsigned alpha); /* he_cmp :460-530 */
void hb_he_cmppt(hb_ctx *ctx, hb_ct *dst, const hb_ct *ct, const hb_pt *pt, const hb_evk *rlk, unsigned iter, unsigned alpha); /* he_cmppt :532-548 */

/* ---- client side on the device: he_enc_pk / he_dec, he-encrypt.c:37-74,107-123 ------------------ */
/* Sampling stays with the caller (the reference's deterministic stream defines every ciphertext): v (sample_zo), e0, e1
 * (sample_error) arrive as one signed word per coefficient, HOST [batch][N]; the plaintexts as HOST [batch][W_L][N]. */
void hb_pk_upload(hb_ctx *ctx, hb_evk *pk, const uint64_t *p0_words, const uint64_t *p1_words);   /* he_pk: HOST [W_L][N] each */
void hb_sk_upload(hb_ctx *ctx, hb_sk *sk, const int64_t *s);                                      /* HOST [N] */
void hb_sk_free(hb_ctx *ctx, hb_sk *sk);
void hb_he_enc_pk(hb_ctx *ctx, hb_ct *ct, const uint64_t *pt_words, double pt_nu, const hb_evk *pk, const int64_t *v,
                  const int64_t *e0, const int64_t *e1);
/* m = smod(c1 sk + c0, q_l): HOST pt_words[batch][W_l][N] (synchronises); the plaintext's nu is ct->nu */
void hb_he_dec(hb_ctx *ctx, uint64_t *pt_words, const hb_ct *ct, const hb_sk *sk);

/* ---- key generation on the device: he_genswk, he-kem.c:74-119 ------------------------------------- */
/* For nkeys keys at once:  p1 = smod(a, P q_L),  p0 = smod(-a sk + e + P sp, P q_L),  both as dimevk transform limbs in the
 * reference's [0,p] representation -- what hb_evk_upload takes.  Sampling stays with the caller: a (sample_uniform below
 * 2^nbits(P q_L)) as HOST words [nkeys][hb_genswk_words][N], e (sample_error) and sp -- the polynomial switched from:
 * poly_rot / poly_conj of sk (he_genrk, he_genck) or sk^2 mod q_L (he_genrlk) -- as HOST int64 [nkeys][N].
 * out_p0, out_p1: HOST [nkeys][dimevk][N].  Synchronises. */
unsigned hb_genswk_words(const hb_ctx *ctx);
void hb_he_genswk(hb_ctx *ctx, unsigned nkeys, const uint64_t *a_words, const int64_t *e, const int64_t *sp, const hb_sk *sk,
                  uint64_t *out_p0, uint64_t *out_p1);

/* ---- wire format (additive: the reference has none).  Layout in hb_wire.cpp. ---------------------------- */
size_t hb_wire_ct_bytes(const hb_ctx *ctx, unsigned l, unsigned batch);
size_t hb_wire_evk_bytes(const hb_ctx *ctx);
size_t hb_ct_serialize(hb_ctx *ctx, const hb_ct *ct, void *buf, size_t cap);          /* bytes written, 0 if cap is too small */
void hb_ct_deserialize(hb_ctx *ctx, hb_ct *ct, const void *buf, size_t len);          /* ct allocated with the message's batch */
size_t hb_evk_serialize(const hb_ctx *ctx, const uint64_t *p0, const uint64_t *p1, void *buf, size_t cap);   /* HOST limbs [dimevk][N] */
void hb_evk_deserialize(hb_ctx *ctx, hb_evk *evk, const void *buf, size_t len);       /* checks, then hb_evk_upload */

/* ---- limb-level entry points (parity tests, NTT microbench) ---------------------------------------- */
/* ntt / invntt (GPQHE/src/ntt.c:37-73) over nlimbs limbs in DEVICE memory; limb j uses prime d0 + j % D.
 * strict != 0 reproduces the reference's [0,p] forward representation bit for bit; strict == 0 returns
 * the same residues canonicalised to [0,p). */
void hb_ntt_forward(hb_ctx *ctx, const uint64_t *in, uint64_t *out, size_t nlimbs, unsigned d0, unsigned D, int strict);
void hb_ntt_inverse(hb_ctx *ctx, const uint64_t *in, uint64_t *out, size_t nlimbs, unsigned d0, unsigned D);
/* host-buffer variant: H2D, transform, D2H, synchronise */
void hb_ntt_forward_host(hb_ctx *ctx, const uint64_t *in, uint64_t *out, size_t nlimbs, unsigned d0, unsigned D, int strict);
/* rns_decompose + ntt (rns.c:37-48): npoly DEVICE polynomials of W words -> out[npoly][D][N] */
void hb_poly_ntt(hb_ctx *ctx, const uint64_t *polys, unsigned W, size_t npoly, unsigned D, uint64_t *out);
/* poly_rns2mpi (poly.c:109-120) with modulus 2^k: limbs[npoly][D][N] -> words[npoly][Wout][N], DEVICE */
void hb_rns2words(hb_ctx *ctx, const uint64_t *limbs, size_t npoly, unsigned D, unsigned k, unsigned Wout, uint64_t *out);
/* poly_rot / poly_conj (poly.c:263-283) on DEVICE polynomials; rot < 0 means conjugation */
void hb_poly_rot(hb_ctx *ctx, const uint64_t *polys, unsigned W, size_t npoly, int rot, uint64_t *out);

/* ---- allocation-free variant --------------------------------------------------------------------------- */
/* No counterpart in the reference.
 * hb_he_mulpt for callers that may not allocate (CUDA graph capture, serving loops, harnesses that forbid
 * cudaMalloc): same inputs, outputs, kernels and results as hb_he_mulpt, but the work buffers of the product sit in
 * one static device array of HB_STATIC_SCRATCH_BYTES that every GPU loading the library holds until exit.  The call
 * makes no cudaMalloc / cudaFree.  It aborts with a message when (3 hw + 2 batch D N) * 8 bytes (D = hb_dim_mulpt,
 * hw = D N rounded up to 32 words) exceed the array, and when per-kernel profiling is on.  The array is shared by
 * every call on a GPU and keeps the last call's contents (each word read is written first in the same call): calls
 * that overlap on the GPU (other streams, other contexts on the device) race; calls on one stream do not. */
#define HB_STATIC_SCRATCH_BYTES (32u << 20)
void hb_he_mulpt_static(hb_ctx *ctx, hb_ct *dst, const hb_ct *src, const hb_pt *pt);

#ifdef __cplusplus
}
#endif
#endif
