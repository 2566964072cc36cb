// image_io.cpp — the image formats on either side of the hot path, host code.
//
//   rtb_jpeg_decode / rtb_jpeg_decode_file / rtb_tex_image_file
//       what `image::open("texture/earthmap.jpg").to_rgb8()` + `ImageTexture::new(data, nx, ny)`
//       do upstream (tests/test.rs:201-205, 489-493; crates image 0.25.1 -> zune-jpeg 0.4.11,
//       Cargo.lock:376-394,1363-1364, not vendored).  Written from the JPEG standard (ITU-T T.81
//       baseline / extended-sequential Huffman, 8 bit) and JFIF: canonical Huffman tables, the
//       integer "slow" inverse DCT with 13-bit constants (Loeffler-Ligtenberg-Moschytz), triangle
//       ("fancy") chroma interpolation for 2x1 / 2x2 subsampling and the 16-bit fixed-point
//       YCbCr -> RGB tables — the arithmetic of the IJG / libjpeg-turbo default decode path, which
//       PIL and OpenCV use, so the result is pinned byte-for-byte against those decoders
//       (tests/test_image_io.py) and against the staged assets/earthmap_1024x512.rgb8.
//   rtb_image_write_png
//       8-bit RGB PNG (zlib "stored" blocks, filter 0) next to the P3/P6 writers of scene.cpp.
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "rtb200.h"
#include "scene.hpp"

namespace {

using rtb::fail;

// ------------------------------------------------------------------------------ JPEG decoder
struct HuffTable {
  bool present = false;
  uint8_t bits[17] = {0};
  uint8_t vals[256] = {0};
  // canonical decode tables (T.81 Annex F.2.2.3): per length the smallest / largest code and
  // the index of its first value; plus an 9-bit direct lookup for short codes
  int32_t mincode[17], maxcode[18], valptr[17];
  uint16_t fast[512];  // (length << 8) | value, 0 = longer than 9 bits
  bool build() {
    int code = 0, k = 0;
    std::memset(fast, 0, sizeof fast);
    for (int l = 1; l <= 16; ++l) {
      valptr[l] = k;
      mincode[l] = code;
      for (int i = 0; i < bits[l]; ++i, ++k, ++code) {
        if (k >= 256) return false;
        if (l <= 9) {
          const int lo = code << (9 - l), n = 1 << (9 - l);
          for (int q = 0; q < n; ++q) fast[lo + q] = (uint16_t)((l << 8) | vals[k]);
        }
      }
      maxcode[l] = bits[l] ? code - 1 : -1;
      if (code > (1 << l)) return false;  // over-subscribed
      code <<= 1;
    }
    maxcode[17] = 0x7fffffff;
    return true;
  }
};

struct Component {
  int id = 0, h = 1, v = 1, tq = 0, td = 0, ta = 0;
  int blocks_w = 0, blocks_h = 0;  // allocated in whole MCUs
  int down_w = 0, down_h = 0;      // ceil(image * samp / max_samp): the real samples
  std::vector<uint8_t> plane;      // blocks_w*8 x blocks_h*8
  int pred = 0;
};

struct BitReader {
  const uint8_t* p;
  const uint8_t* end;
  uint32_t acc = 0;
  int n = 0;
  int marker = 0;  // a marker met inside the entropy-coded segment (0xD0..0xD7, 0xD9, ...)
  void fill() {
    while (n <= 24) {
      int byte = 0;
      if (!marker && p < end) {
        byte = *p++;
        if (byte == 0xFF) {
          while (p < end && *p == 0xFF) ++p;  // fill bytes
          const int m = p < end ? *p++ : 0xD9;
          if (m != 0) {
            marker = m;
            byte = 0;
          }
        }
      }
      acc |= (uint32_t)byte << (24 - n);
      n += 8;
    }
  }
  int peek(int k) { return (int)(acc >> (32 - k)); }
  void skip(int k) { acc <<= k, n -= k; }
  int get(int k) {
    if (k == 0) return 0;
    if (n < k) fill();
    const int v = peek(k);
    skip(k);
    return v;
  }
  void reset() { acc = 0, n = 0, marker = 0; }
};

inline int extend(int v, int t) { return v < (1 << (t - 1)) ? v - (1 << t) + 1 : v; }  // T.81 F.2.2.1

inline int decode_symbol(BitReader& br, const HuffTable& h) {
  if (br.n < 16) br.fill();
  const uint16_t f = h.fast[br.peek(9)];
  if (f) {
    br.skip(f >> 8);
    return f & 255;
  }
  int code = br.peek(9), l = 9;
  br.skip(9);
  for (;;) {
    if (l > 16) return -1;
    if (h.maxcode[l] >= 0 && code <= h.maxcode[l] && code >= h.mincode[l]) break;
    code = (code << 1) | br.get(1);
    ++l;
  }
  return h.vals[h.valptr[l] + code - h.mincode[l]];
}

const uint8_t kZigzag[64] = {0,  1,  8,  16, 9,  2,  3,  10, 17, 24, 32, 25, 18, 11, 4,  5,  12, 19, 26, 33, 40, 48,
                             41, 34, 27, 20, 13, 6,  7,  14, 21, 28, 35, 42, 49, 56, 57, 50, 43, 36, 29, 22, 15, 23,
                             30, 37, 44, 51, 58, 59, 52, 45, 38, 31, 39, 46, 53, 60, 61, 54, 47, 55, 62, 63};

// Output stage of the inverse DCT: level shift by 128 and limit to 0..255, with the 10-bit
// wrap-around the table-driven IJG limiter has for out-of-range (corrupt) input.
inline uint8_t range_limit(int x) {
  const int m = x & 1023;
  return (uint8_t)(m < 128 ? m + 128 : (m < 512 ? 255 : (m < 896 ? 0 : m - 896)));
}

// 8x8 inverse DCT, integer, two passes, 13-bit constants and 2 extra bits kept between the
// passes (the LLM factorisation with 12 multiplies per 1-D transform).
void idct8x8(const int* coef, uint8_t* out, int stride) {
  constexpr int CB = 13, P1 = 2;
  constexpr int F0_298 = 2446, F0_390 = 3196, F0_541 = 4433, F0_765 = 6270, F0_899 = 7373, F1_175 = 9633, F1_501 = 12299,
                F1_847 = 15137, F1_961 = 16069, F2_053 = 16819, F2_562 = 20995, F3_072 = 25172;
  auto descale = [](long x, int n) { return (int)((x + (1L << (n - 1))) >> n); };
  int ws[64];
  for (int pass = 0; pass < 2; ++pass) {
    for (int q = 0; q < 8; ++q) {
      // pass 0 walks columns of coef into ws, pass 1 walks rows of ws into pixels
      const int* in = pass == 0 ? coef + q : ws + 8 * q;
      const int st = pass == 0 ? 8 : 1;
      long z2 = in[2 * st], z3 = in[6 * st];
      long z1 = (z2 + z3) * F0_541;
      long tmp2 = z1 + z3 * (-F1_847), tmp3 = z1 + z2 * F0_765;
      z2 = in[0], z3 = in[4 * st];
      long tmp0 = (z2 + z3) * (1L << CB), tmp1 = (z2 - z3) * (1L << CB);
      const long tmp10 = tmp0 + tmp3, tmp13 = tmp0 - tmp3, tmp11 = tmp1 + tmp2, tmp12 = tmp1 - tmp2;
      tmp0 = in[7 * st], tmp1 = in[5 * st], tmp2 = in[3 * st], tmp3 = in[1 * st];
      z1 = tmp0 + tmp3, z2 = tmp1 + tmp2, z3 = tmp0 + tmp2;
      long z4 = tmp1 + tmp3;
      const long z5 = (z3 + z4) * F1_175;
      tmp0 *= F0_298, tmp1 *= F2_053, tmp2 *= F3_072, tmp3 *= F1_501;
      z1 *= -F0_899, z2 *= -F2_562, z3 *= -F1_961, z4 *= -F0_390;
      z3 += z5, z4 += z5;
      tmp0 += z1 + z3, tmp1 += z2 + z4, tmp2 += z2 + z3, tmp3 += z1 + z4;
      const long o[8] = {tmp10 + tmp3, tmp11 + tmp2, tmp12 + tmp1, tmp13 + tmp0,
                         tmp13 - tmp0, tmp12 - tmp1, tmp11 - tmp2, tmp10 - tmp3};
      if (pass == 0) {
        for (int r = 0; r < 8; ++r) ws[8 * r + q] = descale(o[r], CB - P1);
      } else {
        for (int c = 0; c < 8; ++c) out[q * stride + c] = range_limit(descale(o[c], CB + P1 + 3));
      }
    }
  }
}

struct Decoder {
  const uint8_t* d = nullptr;
  size_t len = 0;
  std::string err;
  uint16_t qt[4][64];
  bool qt_ok[4] = {false, false, false, false};
  HuffTable dc[4], ac[4];
  std::vector<Component> comps;
  int width = 0, height = 0, restart = 0, hmax = 1, vmax = 1;
  bool jfif = false, adobe = false;
  int adobe_transform = 0;

  bool bad(const std::string& m) {
    err = m;
    return false;
  }

  bool decode_block(BitReader& br, Component& c, int* coef) {
    std::memset(coef, 0, 64 * sizeof(int));
    const HuffTable& hd = dc[c.td];
    const HuffTable& ha = ac[c.ta];
    const uint16_t* q = qt[c.tq];
    int t = decode_symbol(br, hd);
    if (t < 0 || t > 15) return bad("bad DC code");
    if (t) c.pred += extend(br.get(t), t);
    coef[0] = c.pred * q[0];
    for (int k = 1; k < 64;) {
      const int rs = decode_symbol(br, ha);
      if (rs < 0) return bad("bad AC code");
      const int r = rs >> 4, s = rs & 15;
      if (s == 0) {
        if (r != 15) break;  // EOB
        k += 16;
        continue;
      }
      k += r;
      if (k > 63) return bad("AC run past the block");
      coef[kZigzag[k]] = extend(br.get(s), s) * q[k];
      ++k;
    }
    return true;
  }

  bool scan(size_t pos, const std::vector<int>& order) {
    // baseline: one interleaved scan over all components, or one scan per component
    BitReader br{d + pos, d + len};
    const bool interleaved = order.size() > 1;
    int mcus_x, mcus_y;
    if (interleaved) {
      mcus_x = (width + 8 * hmax - 1) / (8 * hmax), mcus_y = (height + 8 * vmax - 1) / (8 * vmax);
    } else {
      const Component& c = comps[order[0]];
      mcus_x = (c.down_w + 7) / 8, mcus_y = (c.down_h + 7) / 8;
    }
    for (int ci : order) comps[ci].pred = 0;
    int coef[64];
    int until_restart = restart, next_rst = 0;
    for (int my = 0; my < mcus_y; ++my)
      for (int mx = 0; mx < mcus_x; ++mx) {
        if (restart && until_restart == 0) {
          // byte-align, expect RSTn
          br.n = 0, br.acc = 0;
          if (!br.marker) {  // the marker has not been consumed by look-ahead yet
            while (br.p + 1 < br.end && !(br.p[0] == 0xFF && br.p[1] >= 0xD0 && br.p[1] <= 0xD7)) ++br.p;
            if (br.p + 1 >= br.end) return bad("missing restart marker");
            br.marker = br.p[1];
            br.p += 2;
          }
          if (br.marker != 0xD0 + next_rst) return bad("restart marker out of sequence");
          next_rst = (next_rst + 1) & 7;
          br.reset();
          for (int ci : order) comps[ci].pred = 0;
          until_restart = restart;
        }
        for (int ci : order) {
          Component& c = comps[ci];
          const int bh = interleaved ? c.h : 1, bv = interleaved ? c.v : 1;
          for (int by = 0; by < bv; ++by)
            for (int bx = 0; bx < bh; ++bx) {
              if (!decode_block(br, c, coef)) return false;
              const int X = (mx * bh + bx) * 8, Y = (my * bv + by) * 8;
              idct8x8(coef, c.plane.data() + (size_t)Y * (c.blocks_w * 8) + X, c.blocks_w * 8);
            }
        }
        if (restart) --until_restart;
      }
    return true;
  }

  bool run() {
    if (len < 4 || d[0] != 0xFF || d[1] != 0xD8) return bad("not a JPEG (no SOI)");
    size_t i = 2;
    int scans_done = 0;
    std::vector<bool> seen;
    while (i + 4 <= len) {
      if (d[i] != 0xFF) return bad("marker expected");
      while (i < len && d[i] == 0xFF) ++i;
      if (i >= len) break;
      const int m = d[i++];
      if (m == 0xD9) break;
      if (m == 0x01 || (m >= 0xD0 && m <= 0xD7)) continue;
      if (i + 2 > len) return bad("truncated segment");
      const size_t L = ((size_t)d[i] << 8) | d[i + 1];
      if (L < 2 || i + L > len) return bad("truncated segment");
      const uint8_t* s = d + i + 2;
      const size_t n = L - 2;
      if (m == 0xDB) {  // DQT
        size_t k = 0;
        while (k < n) {
          const int pq = s[k] >> 4, tq = s[k] & 15;
          ++k;
          if (tq > 3 || pq > 1 || k + (pq ? 128 : 64) > n) return bad("bad DQT");
          for (int z = 0; z < 64; ++z) {  // kept in zigzag order; decode_block indexes it by k
            qt[tq][z] = pq ? (uint16_t)((s[k] << 8) | s[k + 1]) : s[k];
            k += pq ? 2 : 1;
          }
          qt_ok[tq] = true;
        }
      } else if (m == 0xC4) {  // DHT
        size_t k = 0;
        while (k < n) {
          if (k + 17 > n) return bad("bad DHT");
          const int tc = s[k] >> 4, th = s[k] & 15;
          if (tc > 1 || th > 3) return bad("bad DHT");
          HuffTable& h = tc ? ac[th] : dc[th];
          int total = 0;
          h.bits[0] = 0;
          for (int l = 1; l <= 16; ++l) total += (h.bits[l] = s[k + l]);
          k += 17;
          if (total > 256 || k + total > n) return bad("bad DHT");
          std::memcpy(h.vals, s + k, total);
          k += total;
          if (!h.build()) return bad("bad DHT (code lengths)");
          h.present = true;
        }
      } else if (m == 0xC0 || m == 0xC1) {  // SOF0 / SOF1
        if (n < 6 || s[0] != 8) return bad("only 8-bit samples are supported");
        height = (s[1] << 8) | s[2], width = (s[3] << 8) | s[4];
        const int nc = s[5];
        if ((nc != 1 && nc != 3) || n < (size_t)(6 + 3 * nc) || width == 0 || height == 0)
          return bad("unsupported frame (components / size)");
        if ((size_t)width * height > ((size_t)1 << 28)) return bad("frame larger than 2^28 pixels");
        comps.assign(nc, Component());
        for (int c = 0; c < nc; ++c) {
          comps[c].id = s[6 + 3 * c], comps[c].h = s[7 + 3 * c] >> 4, comps[c].v = s[7 + 3 * c] & 15;
          comps[c].tq = s[8 + 3 * c];
          if (comps[c].h < 1 || comps[c].h > 2 || comps[c].v < 1 || comps[c].v > 2 || comps[c].tq > 3)
            return bad("unsupported sampling factors");
          hmax = std::max(hmax, comps[c].h), vmax = std::max(vmax, comps[c].v);
        }
        if (nc == 1) comps[0].h = comps[0].v = hmax = vmax = 1;
        const int mcus_x = (width + 8 * hmax - 1) / (8 * hmax), mcus_y = (height + 8 * vmax - 1) / (8 * vmax);
        for (Component& c : comps) {
          if (nc == 3 && &c != &comps[0] && (c.h != 1 || c.v != 1)) return bad("unsupported chroma sampling");
          c.blocks_w = mcus_x * c.h, c.blocks_h = mcus_y * c.v;
          c.down_w = (width * c.h + hmax - 1) / hmax, c.down_h = (height * c.v + vmax - 1) / vmax;
          c.plane.assign((size_t)c.blocks_w * 8 * c.blocks_h * 8, 0);
        }
        if (nc == 3 && (comps[0].h != hmax || comps[0].v != vmax)) return bad("unsupported luma sampling");
        if (nc == 3 && hmax == 1 && vmax == 2) return bad("unsupported sampling (1x2)");
        seen.assign(nc, false);
      } else if (m == 0xC2 || (m >= 0xC3 && m <= 0xCF && m != 0xC4 && m != 0xC8 && m != 0xCC)) {
        return bad("unsupported JPEG process (progressive / lossless / arithmetic)");
      } else if (m == 0xDD) {
        if (n < 2) return bad("bad DRI");
        restart = (s[0] << 8) | s[1];
      } else if (m == 0xE0) {
        if (n >= 5 && std::memcmp(s, "JFIF\0", 5) == 0) jfif = true;
      } else if (m == 0xEE) {
        if (n >= 12 && std::memcmp(s, "Adobe", 5) == 0) adobe = true, adobe_transform = s[11];
      } else if (m == 0xDA) {  // SOS
        if (comps.empty()) return bad("SOS before SOF");
        if (n < 1) return bad("bad SOS");
        const int ns = s[0];
        if (ns < 1 || ns > (int)comps.size() || n < (size_t)(1 + 2 * ns + 3)) return bad("bad SOS");
        std::vector<int> order;
        for (int q = 0; q < ns; ++q) {
          int ci = -1;
          for (size_t c = 0; c < comps.size(); ++c)
            if (comps[c].id == s[1 + 2 * q]) ci = (int)c;
          if (ci < 0) return bad("SOS names an unknown component");
          comps[ci].td = s[2 + 2 * q] >> 4, comps[ci].ta = s[2 + 2 * q] & 15;
          if (comps[ci].td > 3 || comps[ci].ta > 3 || !dc[comps[ci].td].present || !ac[comps[ci].ta].present)
            return bad("SOS names a missing Huffman table");
          if (!qt_ok[comps[ci].tq]) return bad("missing quantisation table");
          order.push_back(ci);
          seen[ci] = true;
        }
        if (ns != 1 && ns != (int)comps.size()) return bad("unsupported partial interleave");
        if (!scan(i + L, order)) return false;
        ++scans_done;
        // continue after the entropy-coded data: the next marker that is not RSTn / stuffing
        size_t k = i + L;
        while (k + 1 < len && !(d[k] == 0xFF && d[k + 1] != 0 && d[k + 1] != 0xFF && !(d[k + 1] >= 0xD0 && d[k + 1] <= 0xD7))) ++k;
        i = k;
        continue;
      }
      i += L;
    }
    if (!scans_done) return bad("no scan");
    for (bool b : seen)
      if (!b) return bad("a component has no scan");
    return true;
  }

  // chroma plane at full resolution: replication is never used, the default decode path
  // interpolates ("fancy" upsampling) with a triangle filter
  std::vector<uint8_t> upsample(const Component& c) const {
    const int W = c.blocks_w * 8;
    const int ow = width, oh = height;
    std::vector<uint8_t> out((size_t)ow * oh);
    if (c.h == hmax && c.v == vmax) {
      for (int y = 0; y < oh; ++y) std::memcpy(&out[(size_t)y * ow], &c.plane[(size_t)y * W], ow);
      return out;
    }
    const int dw = c.down_w, dh = c.down_h;
    if (dw <= 2) {  // too narrow for the triangle filter: plain replication
      for (int y = 0; y < oh; ++y)
        for (int x = 0; x < ow; ++x) out[(size_t)y * ow + x] = c.plane[(size_t)(y * c.v / vmax) * W + x / 2];
      return out;
    }
    std::vector<uint8_t> row(2 * (size_t)dw + 2);
    if (vmax == 1) {  // 2x1: 3/4 nearer + 1/4 further sample, rounding alternately down / up
      for (int y = 0; y < oh; ++y) {
        const uint8_t* in = &c.plane[(size_t)y * W];
        {
          row[0] = in[0];
          row[1] = (uint8_t)((in[0] * 3 + in[1] + 2) >> 2);
          for (int x = 1; x < dw - 1; ++x) {
            row[2 * x] = (uint8_t)((in[x] * 3 + in[x - 1] + 1) >> 2);
            row[2 * x + 1] = (uint8_t)((in[x] * 3 + in[x + 1] + 2) >> 2);
          }
          row[2 * dw - 2] = (uint8_t)((in[dw - 1] * 3 + in[dw - 2] + 1) >> 2);
          row[2 * dw - 1] = in[dw - 1];
        }
        std::memcpy(&out[(size_t)y * ow], row.data(), ow);
      }
      return out;
    }
    // 2x2: vertical 3:1 blend with the nearer neighbour row first (edge rows repeat), then the
    // same 3:1 blend horizontally on the 16x sums, rounding constants 8 / 7
    std::vector<int> sum(dw);
    for (int y = 0; y < oh; ++y) {
      const int cy = y >> 1;
      int ny = (y & 1) ? cy + 1 : cy - 1;
      ny = ny < 0 ? 0 : (ny > dh - 1 ? dh - 1 : ny);
      const uint8_t* in0 = &c.plane[(size_t)cy * W];
      const uint8_t* in1 = &c.plane[(size_t)ny * W];
      for (int x = 0; x < dw; ++x) sum[x] = in0[x] * 3 + in1[x];
      {
        row[0] = (uint8_t)((sum[0] * 4 + 8) >> 4);
        row[1] = (uint8_t)((sum[0] * 3 + sum[1] + 7) >> 4);
        for (int x = 1; x < dw - 1; ++x) {
          row[2 * x] = (uint8_t)((sum[x] * 3 + sum[x - 1] + 8) >> 4);
          row[2 * x + 1] = (uint8_t)((sum[x] * 3 + sum[x + 1] + 7) >> 4);
        }
        row[2 * dw - 2] = (uint8_t)((sum[dw - 1] * 3 + sum[dw - 2] + 8) >> 4);
        row[2 * dw - 1] = (uint8_t)((sum[dw - 1] * 4 + 7) >> 4);
      }
      std::memcpy(&out[(size_t)y * ow], row.data(), ow);
    }
    return out;
  }

  void to_rgb(uint8_t* rgb) const {
    const size_t npix = (size_t)width * height;
    const int W0 = comps[0].blocks_w * 8;
    if (comps.size() == 1) {
      for (int y = 0; y < height; ++y)
        for (int x = 0; x < width; ++x) {
          const uint8_t g = comps[0].plane[(size_t)y * W0 + x];
          uint8_t* o = rgb + 3 * ((size_t)y * width + x);
          o[0] = o[1] = o[2] = g;
        }
      return;
    }
    const std::vector<uint8_t> p1 = upsample(comps[1]), p2 = upsample(comps[2]);
    // colour space of a 3-component file: JFIF says YCbCr; an Adobe marker says what its transform
    // flag says; otherwise component ids 'R','G','B' mean RGB and anything else YCbCr
    bool ycc = true;
    if (!jfif) {
      if (adobe) ycc = adobe_transform != 0;
      else if (comps[0].id == 'R' && comps[1].id == 'G' && comps[2].id == 'B') ycc = false;
    }
    // 16-bit fixed-point conversion tables
    int cr_r[256], cb_b[256];
    long cr_g[256], cb_g[256];
    auto fix = [](double x) { return (long)(x * 65536.0 + 0.5); };
    for (int v = 0; v < 256; ++v) {
      const long x = v - 128;
      cr_r[v] = (int)((fix(1.40200) * x + 32768) >> 16);
      cb_b[v] = (int)((fix(1.77200) * x + 32768) >> 16);
      cr_g[v] = -fix(0.71414) * x;
      cb_g[v] = -fix(0.34414) * x + 32768;
    }
    auto clamp8 = [](int v) { return (uint8_t)(v < 0 ? 0 : (v > 255 ? 255 : v)); };
    for (int y = 0; y < height; ++y)
      for (int x = 0; x < width; ++x) {
        const size_t k = (size_t)y * width + x;
        const int Y = comps[0].plane[(size_t)y * W0 + x], cb = p1[k], cr = p2[k];
        uint8_t* o = rgb + 3 * k;
        if (ycc) {
          o[0] = clamp8(Y + cr_r[cr]);
          o[1] = clamp8(Y + (int)((cb_g[cb] + cr_g[cr]) >> 16));
          o[2] = clamp8(Y + cb_b[cb]);
        } else {
          o[0] = (uint8_t)Y, o[1] = (uint8_t)cb, o[2] = (uint8_t)cr;
        }
      }
    (void)npix;
  }
};

bool read_file(const char* path, std::vector<uint8_t>& out) {
  FILE* f = std::fopen(path, "rb");
  if (!f) return false;
  uint8_t buf[65536];
  size_t n;
  while ((n = std::fread(buf, 1, sizeof buf, f)) > 0) out.insert(out.end(), buf, buf + n);
  std::fclose(f);
  return true;
}

// ------------------------------------------------------------------------------ PNG writer
uint32_t crc32_update(uint32_t c, const uint8_t* p, size_t n) {
  static uint32_t table[256];
  static bool init = false;
  if (!init) {
    for (uint32_t i = 0; i < 256; ++i) {
      uint32_t v = i;
      for (int k = 0; k < 8; ++k) v = (v & 1) ? 0xEDB88320u ^ (v >> 1) : v >> 1;
      table[i] = v;
    }
    init = true;
  }
  for (size_t i = 0; i < n; ++i) c = table[(c ^ p[i]) & 255] ^ (c >> 8);
  return c;
}
void put32(std::vector<uint8_t>& v, uint32_t x) {
  v.push_back((uint8_t)(x >> 24)), v.push_back((uint8_t)(x >> 16)), v.push_back((uint8_t)(x >> 8)), v.push_back((uint8_t)x);
}
void chunk(std::vector<uint8_t>& png, const char* type, const std::vector<uint8_t>& data) {
  put32(png, (uint32_t)data.size());
  const size_t at = png.size();
  png.insert(png.end(), type, type + 4);
  png.insert(png.end(), data.begin(), data.end());
  put32(png, crc32_update(0xFFFFFFFFu, png.data() + at, png.size() - at) ^ 0xFFFFFFFFu);
}

}  // namespace

extern "C" {

int rtb_jpeg_decode(const uint8_t* data, size_t len, uint8_t** rgb8_out, uint32_t* nx, uint32_t* ny) {
  if (!data || !rgb8_out || !nx || !ny) return fail(RTB_ERR_INVALID, "jpeg_decode: null argument");
  *rgb8_out = nullptr, *nx = 0, *ny = 0;
  Decoder dec;
  dec.d = data, dec.len = len;
  if (!dec.run()) return fail(RTB_ERR_INVALID, "jpeg_decode: " + dec.err);
  uint8_t* rgb = (uint8_t*)std::malloc((size_t)dec.width * dec.height * 3);
  if (!rgb) return fail(RTB_ERR_INVALID, "jpeg_decode: out of memory");
  dec.to_rgb(rgb);
  *rgb8_out = rgb, *nx = (uint32_t)dec.width, *ny = (uint32_t)dec.height;
  return RTB_OK;
}

int rtb_jpeg_decode_file(const char* path, uint8_t** rgb8_out, uint32_t* nx, uint32_t* ny) {
  std::vector<uint8_t> bytes;
  if (!path || !read_file(path, bytes)) return fail(RTB_ERR_IO, std::string("jpeg_decode_file: cannot read ") + (path ? path : "(null)"));
  return rtb_jpeg_decode(bytes.data(), bytes.size(), rgb8_out, nx, ny);
}

void rtb_image_free(uint8_t* rgb8) { std::free(rgb8); }

int rtb_tex_image_file(rtb_scene* s, const char* path) {
  // image::open(path).expect(..).to_rgb8(); ImageTexture::new(data, nx, ny)  (tests/test.rs:201-205)
  uint8_t* rgb = nullptr;
  uint32_t nx = 0, ny = 0;
  const int rc = rtb_jpeg_decode_file(path, &rgb, &nx, &ny);
  if (rc != RTB_OK) return rc;
  const int h = rtb_tex_image(s, rgb, nx, ny);
  std::free(rgb);
  return h;
}

int rtb_image_write_png(const char* path, const uint8_t* rgb8, int nx, int ny) {
  if (!path || !rgb8 || nx <= 0 || ny <= 0) return fail(RTB_ERR_INVALID, "image_write_png: bad argument");
  std::vector<uint8_t> png = {0x89, 'P', 'N', 'G', 0x0D, 0x0A, 0x1A, 0x0A};
  std::vector<uint8_t> ihdr;
  put32(ihdr, (uint32_t)nx), put32(ihdr, (uint32_t)ny);
  ihdr.insert(ihdr.end(), {8, 2, 0, 0, 0});  // 8 bit, truecolour, deflate, adaptive filtering, no interlace
  chunk(png, "IHDR", ihdr);
  // zlib stream of stored blocks over (filter byte 0 + row) x ny
  const size_t row = 1 + 3 * (size_t)nx, raw_n = row * (size_t)ny;
  std::vector<uint8_t> z;
  z.reserve(raw_n + raw_n / 65535 * 5 + 16);
  z.push_back(0x78), z.push_back(0x01);
  uint32_t a = 1, b = 0;  // adler32
  std::vector<uint8_t> raw(raw_n);
  for (int y = 0; y < ny; ++y) {
    raw[y * row] = 0;
    std::memcpy(&raw[y * row + 1], rgb8 + 3 * (size_t)nx * y, 3 * (size_t)nx);
  }
  for (size_t off = 0; off < raw_n;) {
    const size_t n = std::min<size_t>(65535, raw_n - off);
    z.push_back(off + n == raw_n ? 1 : 0);
    z.push_back((uint8_t)(n & 255)), z.push_back((uint8_t)(n >> 8));
    z.push_back((uint8_t)(~n & 255)), z.push_back((uint8_t)((~n >> 8) & 255));
    z.insert(z.end(), raw.begin() + off, raw.begin() + off + n);
    off += n;
  }
  for (size_t i = 0; i < raw_n;) {
    const size_t n = std::min<size_t>(5552, raw_n - i);
    for (size_t k = 0; k < n; ++k) a += raw[i + k], b += a;
    a %= 65521u, b %= 65521u;
    i += n;
  }
  put32(z, (b << 16) | a);
  chunk(png, "IDAT", z);
  chunk(png, "IEND", {});
  FILE* f = std::fopen(path, "wb");
  if (!f) return fail(RTB_ERR_IO, std::string("image_write_png: cannot create ") + path);
  const size_t w = std::fwrite(png.data(), 1, png.size(), f);
  if (std::fclose(f) != 0 || w != png.size()) return fail(RTB_ERR_IO, std::string("image_write_png: short write to ") + path);
  return RTB_OK;
}

}  // extern "C"
