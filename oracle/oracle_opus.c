/* oracle_opus.c -- ORACLE (test infrastructure): plain-C restatement of the reference's Opus
 * decode layer for CELT-only packets -- TOC / frame packing parse (opus-fix/src/opus.c:148-343),
 * opus_decode_native / opus_decode_frame (src/opus_decoder.c:200-709) incl. PLC chunking and the
 * decoder gain, and the multistream wrapper (src/opus_multistream_decoder.c:155-302).
 * SILK / hybrid packets are outside the path built so far and return OPUS_UNIMPLEMENTED (-5). */
#include <stdlib.h>
#include <string.h>
#include "oracle_celt.h"

#define MODE_SILK 1000
#define MODE_HYBRID 1001
#define MODE_CELT 1002

struct OracleOpus {
   int Fs, channels, decode_gain;
   int stream_channels, bandwidth, mode, prev_mode, frame_size, prev_redundancy, last_packet_duration;
   uint32_t rangeFinal;
   struct OracleCelt *celt;
};

static int imin_(int a, int b) { return a < b ? a : b; }

static int toc_mode(int t) { return (t & 0x80) ? MODE_CELT : ((t & 0x60) == 0x60) ? MODE_HYBRID : MODE_SILK; }   /* opus_decoder.c:184 */
static int toc_bandwidth(int t) {                                                                               /* :921-937 */
   if (t & 0x80) { int bw = 1102 + ((t >> 5) & 3); return bw == 1102 ? 1101 : bw; }
   if ((t & 0x60) == 0x60) return (t & 0x10) ? 1105 : 1104;
   return 1101 + ((t >> 5) & 3);
}
static int toc_samples(int t, int Fs) {                                                                          /* opus.c:169-188 */
   if (t & 0x80) return (Fs << ((t >> 3) & 3)) / 400;
   if ((t & 0x60) == 0x60) return (t & 8) ? Fs / 50 : Fs / 100;
   { int a = (t >> 3) & 3; return a == 3 ? Fs * 60 / 1000 : (Fs << a) / 100; }
}
static int parse_size(const uint8_t *d, int len, int *size) {                                                    /* opus.c:148-167 */
   if (len < 1) { *size = -1; return -1; }
   if (d[0] < 252) { *size = d[0]; return 1; }
   if (len < 2) { *size = -1; return -1; }
   *size = 4 * d[1] + d[0]; return 2;
}
/* opus_packet_parse_impl (opus.c:190-343) */
static int parse_packet(const uint8_t *data, int len, int self_delimited, int *out_toc, int16_t size[48],
                        int *payload_offset, int *packet_offset) {
   const uint8_t *d0 = data; int i, bytes, count, cbr = 0, toc, framesize, last_size, pad = 0, sz;
   if (len < 0) return -1;
   if (len == 0) return -4;
   framesize = toc_samples(data[0], 48000);
   toc = *data++; len--; last_size = len;
   switch (toc & 3) {
      case 0: count = 1; break;
      case 1: count = 2; cbr = 1;
         if (!self_delimited) { if (len & 1) return -4; last_size = len / 2; size[0] = (int16_t)last_size; }
         break;
      case 2: count = 2; bytes = parse_size(data, len, &sz); size[0] = (int16_t)sz; len -= bytes;
         if (sz < 0 || sz > len) return -4;
         data += bytes; last_size = len - sz; break;
      default: {
         int ch;
         if (len < 1) return -4;
         ch = *data++; count = ch & 0x3F;
         if (count <= 0 || framesize * count > 5760) return -4;
         len--;
         if (ch & 0x40) { int p; do { int tmp; if (len <= 0) return -4; p = *data++; len--; tmp = p == 255 ? 254 : p; len -= tmp; pad += tmp; } while (p == 255); }
         if (len < 0) return -4;
         cbr = !(ch & 0x80);
         if (!cbr) {
            last_size = len;
            for (i = 0; i < count - 1; i++) {
               bytes = parse_size(data, len, &sz); size[i] = (int16_t)sz; len -= bytes;
               if (sz < 0 || sz > len) return -4;
               data += bytes; last_size -= bytes + sz;
            }
            if (last_size < 0) return -4;
         } else if (!self_delimited) {
            last_size = len / count;
            if (last_size * count != len) return -4;
            for (i = 0; i < count - 1; i++) size[i] = (int16_t)last_size;
         }
      }
   }
   if (self_delimited) {
      bytes = parse_size(data, len, &sz); size[count - 1] = (int16_t)sz; len -= bytes;
      if (sz < 0 || sz > len) return -4;
      data += bytes;
      if (cbr) { if (sz * count > len) return -4; for (i = 0; i < count - 1; i++) size[i] = (int16_t)sz; }
      else if (bytes + sz > last_size) return -4;
   } else {
      if (last_size > 1275) return -4;
      size[count - 1] = (int16_t)last_size;
   }
   if (payload_offset) *payload_offset = (int)(data - d0);
   for (i = 0; i < count; i++) data += size[i];
   if (packet_offset) *packet_offset = pad + (int)(data - d0);
   if (out_toc) *out_toc = toc;
   return count;
}
static int endband(int bw) { return bw == 1101 ? 13 : (bw == 1102 || bw == 1103) ? 17 : bw == 1104 ? 19 : 21; }   /* opus_decoder.c:432-449 */

/* opus_decode_frame (opus_decoder.c:200-596), CELT-only branch */
static int decode_frame(struct OracleOpus *st, const uint8_t *data, int len, int16_t *pcm, int frame_size) {
   int F20 = st->Fs / 50, F10 = F20 >> 1, F5 = F10 >> 1, F2_5 = F5 >> 1, audiosize, mode, i, ret;
   uint32_t rng = 0;
   if (frame_size < F2_5) return -2;
   frame_size = imin_(frame_size, st->Fs / 25 * 3);
   if (len <= 1) { data = NULL; frame_size = imin_(frame_size, st->frame_size); }
   if (data != NULL) { audiosize = st->frame_size; mode = st->mode; }
   else {
      audiosize = frame_size; mode = st->prev_mode;
      if (mode == 0) { for (i = 0; i < audiosize * st->channels; i++) pcm[i] = 0; return audiosize; }
      if (audiosize > F20) {
         do { int r = decode_frame(st, NULL, 0, pcm, imin_(audiosize, F20)); if (r < 0) return r; pcm += r * st->channels; audiosize -= r; } while (audiosize > 0);
         return frame_size;
      } else if (audiosize < F20) {
         if (audiosize > F10) audiosize = F10;
         else if (mode != MODE_SILK && audiosize > F5 && audiosize < F10) audiosize = F5;
      }
   }
   if (audiosize > frame_size) return -1;
   frame_size = audiosize;
   if (mode != MODE_CELT) return -5;
   oracle_celt_set_bands(st->celt, 0, endband(st->bandwidth));
   ret = oracle_celt_decode(st->celt, data, len, pcm, imin_(F20, frame_size) * (48000 / st->Fs), st->stream_channels, 0, &rng);
   if (st->decode_gain) {                                                                                        /* :567-577 */
      extern int32_t oracle_gain_q16(int decode_gain);
      int32_t gain = oracle_gain_q16(st->decode_gain);
      for (i = 0; i < frame_size * st->channels; i++) {
         int32_t x = (int32_t)(((int64_t)pcm[i] * gain + 32768) >> 16);
         pcm[i] = (int16_t)(x > 32767 ? 32767 : x < -32767 ? -32767 : x);
      }
   }
   st->rangeFinal = len <= 1 ? 0 : rng;
   st->prev_mode = mode; st->prev_redundancy = 0;
   return ret < 0 ? ret : audiosize;
}
/* opus_decode_native (opus_decoder.c:598-709) */
static int decode_native(struct OracleOpus *st, const uint8_t *data, int len, int16_t *pcm, int frame_size, int fec,
                         int self_delimited, int *packet_offset) {
   int i, nb, count, offset, toc, pfs, pmode, pbw, pch; int16_t size[48];
   if (fec < 0 || fec > 1) return -1;
   if ((fec || len == 0 || data == NULL) && frame_size % (st->Fs / 400) != 0) return -1;
   if (len == 0 || data == NULL) {
      int cnt = 0;
      do { int r = decode_frame(st, NULL, 0, pcm + cnt * st->channels, frame_size - cnt); if (r < 0) return r; cnt += r; } while (cnt < frame_size);
      st->last_packet_duration = cnt;
      return cnt;
   } else if (len < 0) return -1;
   pmode = toc_mode(data[0]); pbw = toc_bandwidth(data[0]); pfs = toc_samples(data[0], st->Fs); pch = (data[0] & 4) ? 2 : 1;
   count = parse_packet(data, len, self_delimited, &toc, size, &offset, packet_offset);
   if (count < 0) return count;
   data += offset;
   if (fec) {
      if (frame_size < pfs || pmode == MODE_CELT || st->mode == MODE_CELT) return decode_native(st, NULL, 0, pcm, frame_size, 0, 0, NULL);
      return -5;
   }
   if (count * pfs > frame_size) return -2;
   if (pmode != MODE_CELT) return -5;
   st->mode = pmode; st->bandwidth = pbw; st->frame_size = pfs; st->stream_channels = pch;
   nb = 0;
   for (i = 0; i < count; i++) {
      int r = decode_frame(st, data, size[i], pcm + nb * st->channels, frame_size - nb);
      if (r < 0) return r;
      data += size[i]; nb += r;
   }
   st->last_packet_duration = nb;
   return nb;
}

struct OracleOpus *oracle_opus_create(int Fs, int channels) {
   struct OracleOpus *st = (struct OracleOpus *)calloc(1, sizeof(*st));
   st->Fs = Fs; st->channels = st->stream_channels = channels; st->frame_size = Fs / 400;
   st->celt = oracle_celt_create(channels, 48000 / Fs);
   return st;
}
void oracle_opus_destroy(struct OracleOpus *st) { if (st) { oracle_celt_destroy(st->celt); free(st); } }
int oracle_opus_decode(struct OracleOpus *st, const uint8_t *data, int len, int16_t *pcm, int frame_size, int fec) {
   if (frame_size <= 0) return -1;
   return decode_native(st, data, len, pcm, frame_size, fec, 0, NULL);
}
uint32_t oracle_opus_final_range(const struct OracleOpus *st) { return st->rangeFinal; }
void oracle_opus_set_gain(struct OracleOpus *st, int gain) { st->decode_gain = gain; }

/* ---------- multistream: src/opus_multistream_decoder.c ---------- */
struct OracleMS { int Fs, channels, streams, coupled; uint8_t mapping[256]; struct OracleOpus *dec[16]; };

struct OracleMS *oracle_ms_create(int Fs, int channels, int streams, int coupled, const uint8_t *mapping) {
   struct OracleMS *m = (struct OracleMS *)calloc(1, sizeof(*m)); int s;
   m->Fs = Fs; m->channels = channels; m->streams = streams; m->coupled = coupled;
   memcpy(m->mapping, mapping, channels);
   for (s = 0; s < streams; s++) m->dec[s] = oracle_opus_create(Fs, s < coupled ? 2 : 1);
   return m;
}
void oracle_ms_destroy(struct OracleMS *m) { int s; for (s = 0; s < m->streams; s++) oracle_opus_destroy(m->dec[s]); free(m); }
uint32_t oracle_ms_final_range(const struct OracleMS *m) { uint32_t v = 0; int s; for (s = 0; s < m->streams; s++) v ^= m->dec[s]->rangeFinal; return v; }
static int ms_validate(const uint8_t *data, int len, int nb_streams, int Fs) {                                   /* :155-183 */
   int s, samples = 0; int16_t size[48];
   for (s = 0; s < nb_streams; s++) {
      int toc, po, count, cnt, tmp;
      if (len <= 0) return -4;
      count = parse_packet(data, len, s != nb_streams - 1, &toc, size, NULL, &po);
      if (count < 0) return count;
      if (po < 1) return -1;
      cnt = (data[0] & 3) == 0 ? 1 : (data[0] & 3) != 3 ? 2 : po < 2 ? -4 : (data[1] & 0x3F);
      if (cnt < 0) return cnt;
      tmp = cnt * toc_samples(data[0], Fs);
      if (tmp * 25 > Fs * 3) return -4;
      if (s != 0 && samples != tmp) return -4;
      samples = tmp; data += po; len -= po;
   }
   return samples;
}
int oracle_ms_decode(struct OracleMS *m, const uint8_t *data, int len, int16_t *pcm, int frame_size, int fec) {     /* :184-302 */
   static int16_t buf[2 * 5760]; int s, c, t, do_plc = 0;
   if (frame_size <= 0) return -1;
   frame_size = imin_(frame_size, m->Fs / 25 * 3);
   if (data == NULL) len = 0;
   if (len == 0) do_plc = 1;
   if (len < 0) return -1;
   if (!do_plc && len < 2 * m->streams - 1) return -4;
   if (!do_plc) { int r = ms_validate(data, len, m->streams, m->Fs); if (r < 0) return r; if (r > frame_size) return -2; }
   for (s = 0; s < m->streams; s++) {
      int po = 0, ret;
      if (!do_plc && len <= 0) return -3;
      ret = decode_native(m->dec[s], data, len, buf, frame_size, fec, s != m->streams - 1, &po);
      data += po; len -= po;
      if (ret <= 0) return ret;
      frame_size = ret;
      for (c = 0; c < m->channels; c++) {
         int mp = m->mapping[c];
         if (s < m->coupled) { if (mp == 2 * s) for (t = 0; t < frame_size; t++) pcm[t * m->channels + c] = buf[2 * t];
                               if (mp == 2 * s + 1) for (t = 0; t < frame_size; t++) pcm[t * m->channels + c] = buf[2 * t + 1]; }
         else if (mp == s + m->coupled) for (t = 0; t < frame_size; t++) pcm[t * m->channels + c] = buf[t];
      }
   }
   for (c = 0; c < m->channels; c++) if (m->mapping[c] == 255) for (t = 0; t < frame_size; t++) pcm[t * m->channels + c] = 0;
   return frame_size;
}
