// A wire format for ciphertexts and evaluation keys.  The reference has none -- he_ct_t / he_evk_t
// only live in one process (GPQHE/src/precomp.c:455-463 sizes them, nothing writes them) -- so this is additive: what a
// client process and the GPU server exchange.  Little-endian, self-describing, checksummed:
//
//   offset  0  char[4]  "HBCT" | "HBEK"
//           4  u32      version (1)
//           8  u32      logn, logq, logDelta, slots          (the four hectx_init parameters: both ends must agree)
//          24  u32      level l (ct) | dimevk (evk)
//          28  u32      batch (ct)   | 0
//          32  u32      W = words per coefficient (ct) | 1
//          36  u32      reserved (0)
//          40  f64      nu, B                                (he_ct_t bookkeeping, gpqhe.h:84-91; 0 for keys)
//          56  u64      payload bytes
//          64  u64      FNV-1a 64 of the payload
//          72  payload  ct: u64 words[batch][2][W][N] (two's complement, planar)   evk: u64 limbs[2][dimevk][N]
//
// Pure host code over the C ABI: a ciphertext is downloaded / uploaded with hb_ct_download / hb_ct_upload.
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>
#include "../../include/hectr_b200.h"

namespace {

constexpr size_t HEADER = 72;
constexpr uint32_t VERSION = 1;

struct Header {
  char magic[4];
  uint32_t version, logn, logq, logDelta, slots, a, b, c, reserved;
  double nu, B;
  uint64_t payload, checksum;
};

uint64_t fnv1a(const uint8_t *p, size_t n) {
  uint64_t h = 0xcbf29ce484222325ull;
  for (size_t i = 0; i < n; i++) { h ^= p[i]; h *= 0x100000001b3ull; }
  return h;
}

void put_header(uint8_t *out, const Header &h) {
  memcpy(out, h.magic, 4);
  const uint32_t u[9] = {h.version, h.logn, h.logq, h.logDelta, h.slots, h.a, h.b, h.c, h.reserved};
  memcpy(out + 4, u, sizeof u);
  memcpy(out + 40, &h.nu, 8);
  memcpy(out + 48, &h.B, 8);
  memcpy(out + 56, &h.payload, 8);
  memcpy(out + 64, &h.checksum, 8);
}

Header get_header(const uint8_t *in) {
  Header h;
  memcpy(h.magic, in, 4);
  uint32_t u[9];
  memcpy(u, in + 4, sizeof u);
  h.version = u[0]; h.logn = u[1]; h.logq = u[2]; h.logDelta = u[3]; h.slots = u[4]; h.a = u[5]; h.b = u[6]; h.c = u[7]; h.reserved = u[8];
  memcpy(&h.nu, in + 40, 8);
  memcpy(&h.B, in + 48, 8);
  memcpy(&h.payload, in + 56, 8);
  memcpy(&h.checksum, in + 64, 8);
  return h;
}

[[noreturn]] void wire_die(const char *msg) {
  fprintf(stderr, "\033[1m\033[31merror:\033[0m \033[1mhectr_b200\033[0m. Invalid argument. %s\n", msg);
  abort();
}

// 0 = fine, otherwise why not
const char *check(const Header &h, const char magic[4], const hb_info &i, size_t len) {
  if (memcmp(h.magic, magic, 4) != 0) return "wire: wrong magic";
  if (h.version != VERSION) return "wire: unknown version";
  if (h.logn != i.logn || h.logq != i.logq || h.logDelta != i.logDelta || h.slots != i.slots) return "wire: parameters differ from this context";
  if (len < HEADER || h.payload != len - HEADER) return "wire: truncated or oversized message";
  return nullptr;
}

}  // namespace

extern "C" {

size_t hb_wire_ct_bytes(const hb_ctx *ctx, unsigned l, unsigned batch) {
  hb_info i;
  hb_ctx_info(ctx, &i);
  return HEADER + (size_t)batch * 2 * hb_ctx_words(ctx, l) * i.n * 8;
}

size_t hb_wire_evk_bytes(const hb_ctx *ctx) {
  hb_info i;
  hb_ctx_info(ctx, &i);
  return HEADER + (size_t)2 * i.dimevk * i.n * 8;
}

// returns the bytes written, 0 when cap is too small (nothing written)
size_t hb_ct_serialize(hb_ctx *ctx, const hb_ct *ct, void *buf, size_t cap) {
  const size_t need = hb_wire_ct_bytes(ctx, ct->l, ct->batch);
  if (cap < need) return 0;
  hb_info i;
  hb_ctx_info(ctx, &i);
  uint8_t *out = (uint8_t *)buf;
  std::vector<uint64_t> words((need - HEADER) / 8);
  hb_ct_download(ctx, ct, words.data());                      // synchronises
  memcpy(out + HEADER, words.data(), need - HEADER);
  Header h{{'H', 'B', 'C', 'T'}, VERSION, i.logn, i.logq, i.logDelta, i.slots, ct->l, ct->batch, hb_ctx_words(ctx, ct->l), 0,
           ct->nu, ct->B, need - HEADER, fnv1a(out + HEADER, need - HEADER)};
  put_header(out, h);
  return need;
}

// ct must be allocated (hb_ct_alloc) with the message's batch; aborts with the reference's error style on any mismatch
void hb_ct_deserialize(hb_ctx *ctx, hb_ct *ct, const void *buf, size_t len) {
  if (len < HEADER) wire_die("wire: truncated or oversized message");
  const uint8_t *in = (const uint8_t *)buf;
  const Header h = get_header(in);
  hb_info i;
  hb_ctx_info(ctx, &i);
  if (const char *why = check(h, "HBCT", i, len)) wire_die(why);
  if (h.a > i.L) wire_die("wire: level outside the chain");
  if (h.b != ct->batch) wire_die("wire: batch differs from the allocated ciphertext");
  if (h.c != hb_ctx_words(ctx, h.a) || h.payload != (uint64_t)h.b * 2 * h.c * i.n * 8) wire_die("wire: payload size does not match level and batch");
  if (fnv1a(in + HEADER, h.payload) != h.checksum) wire_die("wire: checksum mismatch");
  std::vector<uint64_t> words(h.payload / 8);
  memcpy(words.data(), in + HEADER, h.payload);
  hb_ct_upload(ctx, ct, words.data(), h.a, h.nu, h.B);
}

// keys travel as the reference holds them: p0, p1 = u64[dimevk][N] transform limbs (he-kem.c:103-110), HOST arrays
size_t hb_evk_serialize(const hb_ctx *ctx, const uint64_t *p0, const uint64_t *p1, void *buf, size_t cap) {
  const size_t need = hb_wire_evk_bytes(ctx);
  if (cap < need) return 0;
  hb_info i;
  hb_ctx_info(ctx, &i);
  uint8_t *out = (uint8_t *)buf;
  const size_t half = (need - HEADER) / 2;
  memcpy(out + HEADER, p0, half);
  memcpy(out + HEADER + half, p1, half);
  Header h{{'H', 'B', 'E', 'K'}, VERSION, i.logn, i.logq, i.logDelta, i.slots, i.dimevk, 0, 1, 0, 0.0, 0.0, need - HEADER,
           fnv1a(out + HEADER, need - HEADER)};
  put_header(out, h);
  return need;
}

// parses, checks and uploads (hb_evk_upload)
void hb_evk_deserialize(hb_ctx *ctx, hb_evk *evk, const void *buf, size_t len) {
  if (len < HEADER) wire_die("wire: truncated or oversized message");
  const uint8_t *in = (const uint8_t *)buf;
  const Header h = get_header(in);
  hb_info i;
  hb_ctx_info(ctx, &i);
  if (const char *why = check(h, "HBEK", i, len)) wire_die(why);
  if (h.a != i.dimevk || h.payload != (uint64_t)2 * i.dimevk * i.n * 8) wire_die("wire: key size does not match this context");
  if (fnv1a(in + HEADER, h.payload) != h.checksum) wire_die("wire: checksum mismatch");
  std::vector<uint64_t> limbs(h.payload / 8);
  memcpy(limbs.data(), in + HEADER, h.payload);
  hb_evk_upload(ctx, evk, limbs.data(), limbs.data() + limbs.size() / 2);
}

}  // extern "C"
