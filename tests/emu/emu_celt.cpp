// tests/emu/emu_celt.cpp -- TEST-ONLY build of the kernel bodies as plain C++ (CB_HOST_EMU):
// every lane loop becomes an ordinary loop, every warp fold an identity.  It lets the CPU-side
// tests (-m "not gpu") check the codec logic of csrc/*.h against the oracle without a GPU.
// It is NOT part of the product: libopus_b200.so never links it and has no CPU path.
#define CB_HOST_EMU 1
#include <stdlib.h>
#include <string.h>
#include "../../concentus_b200/csrc/cb_decode_pass.h"
#include "../../concentus_b200/csrc/cb_silk_pass.h"
#include "../../concentus_b200/csrc/cb_opus_general.h"

extern "C" {

// path counters of the band rebuild (cb_celt_bands3.h: CB_COV): copy out and clear
int emu_cov_read(long* out, int n) {
  for (int i = 0; i < n && i < CB_COV_COUNT; i++) { out[i] = cb_cov[i]; cb_cov[i] = 0; }
  return CB_COV_COUNT;
}

// ---- CELT layer only (payload without TOC) ----
void* emu_celt_create(int channels, int downsample) {
  CeltDecState* s = (CeltDecState*)malloc(sizeof(CeltDecState));
  celt_state_reset(s, channels, downsample);
  return s;
}
void emu_celt_destroy(void* s) { free(s); }

int emu_celt_decode(void* state, const unsigned char* payload, int len, int LM, int C, int start, int end,
                    short* pcm, unsigned* final_range) {
  CeltDecState* st = (CeltDecState*)state;
  static EntWork ew;
  static SynWork sw;
  static FrameSide side;
  static FrameTape tape;
  EntFrameCfg cfg = {LM, C, start, end, 0};
  int rc = celt_entropy_decode(ew, st, &side, &tape, payload, len, cfg, final_range);
  if (rc < 0) return rc;
  for (int c = 0; c < st->channels; c++) celt_synth_channel(sw, st, &side, c, pcm);
  return (120 << LM) / st->downsample;
}

// ---- Opus layer: the same plan/pass sequence the CUDA launchers run, for one stream ----
void* emu_opus_create(int Fs, int channels) {
  OpusDecState* s = (OpusDecState*)malloc(sizeof(OpusDecState));
  s->decode_gain = 0;
  opus_state_reset(s, Fs, channels, 0);
  return s;
}
void emu_opus_destroy(void* s) { free(s); }
unsigned emu_opus_final_range(void* s) { return ((OpusDecState*)s)->rangeFinal; }
void emu_opus_set_gain(void* s, int g) { ((OpusDecState*)s)->decode_gain = g; }
int emu_opus_last_packet_duration(void* s) { return ((OpusDecState*)s)->last_packet_duration; }

// max_passes: what a device-resident entry point would be given as max_frames (the host entry points count the
// passes from the packets and never trip the check)
int emu_opus_decode_capped(void* state, const unsigned char* data, int len, short* pcm, int frame_size, int fec,
                           int max_passes) {
  OpusDecState* st = (OpusDecState*)state;
  static EntWork ew;
  static SynWork sw;
  static FrameSide side;
  static PlcWork pw;
  static unsigned char pkt[1600];
  static SilkCallWork cw;
  static GenWork gw;
  static FrameTape tape;
  StreamPlan pl;
  i32 ret = 0;
  opus_plan_stream(st, &pl, data, data ? 0 : -1, data ? len : 0, frame_size, fec, 0, nullptr, max_passes);
  ret = pl.ret;
  for (int pass = 0; pass < pl.nframes; pass++) {
    decode_silk_pass(st, &pl, &side, data, pcm, &ret, pass, cw);
    side.route = pl.route;
    decode_entropy_pass<false>(ew, pkt, 256, st, &pl, &side, &tape, data, &ret, pass);
    decode_entropy_pass<true>(ew, pkt, 256, st, &pl, &side, &tape, data, &ret, pass);
    for (int c = 0; c < st->channels; c++) decode_synth_pass(sw, st, &pl, &side, pcm, c, pass);
    decode_plc_pass(pw, st, &pl, &side, pcm, pass);
    decode_general_pass(gw, st, &pl, &side, &tape, data, pcm, &ret, pass, cw);
  }
  return ret;
}

int emu_opus_decode(void* state, const unsigned char* data, int len, short* pcm, int frame_size, int fec) {
  return emu_opus_decode_capped(state, data, len, pcm, frame_size, fec, CB_MAX_FRAMES);
}

// ---- multistream: plan + passes + channel copy-out, as the CUDA launch sequence does ----
struct EmuMS {
  MSLayout L;
  OpusDecState* st[16];
  StreamPlan plans[16];
};
void* emu_ms_create(int Fs, int channels, int streams, int coupled, const unsigned char* mapping) {
  EmuMS* m = (EmuMS*)calloc(1, sizeof(EmuMS));
  m->L.nb_channels = channels; m->L.nb_streams = streams; m->L.nb_coupled = coupled; m->L.first = 0; m->L.Fs = Fs;
  memcpy(m->L.mapping, mapping, channels);
  for (int s = 0; s < streams; s++) m->st[s] = (OpusDecState*)emu_opus_create(Fs, s < coupled ? 2 : 1);
  return m;
}
void emu_ms_destroy(void* p) {
  EmuMS* m = (EmuMS*)p;
  for (int s = 0; s < m->L.nb_streams; s++) free(m->st[s]);
  free(m);
}
unsigned emu_ms_final_range(void* p) {
  EmuMS* m = (EmuMS*)p;
  unsigned v = 0;
  for (int s = 0; s < m->L.nb_streams; s++) v ^= m->st[s]->rangeFinal;
  return v;
}
int emu_ms_decode(void* p, const unsigned char* data, int len, short* pcm, int frame_size, int fec) {
  EmuMS* m = (EmuMS*)p;
  static EntWork ew;
  static SynWork sw;
  static FrameSide side;
  static PlcWork pw;
  static unsigned char pkt[1600];
  static short epcm[16][5760 * 2];
  static SilkCallWork cw;
  static GenWork gw;
  static FrameTape tape;
  const int S = m->L.nb_streams;
  int ret = opus_ms_plan(m->st, m->plans, &m->L, data, data ? 0 : -1, data ? len : 0, frame_size, fec);
  for (int s = 0; s < S; s++) {
    StreamPlan* pl = &m->plans[s];
    i32 r = pl->ret;
    for (int pass = 0; pass < pl->nframes; pass++) {
      decode_silk_pass(m->st[s], pl, &side, data, epcm[s], &r, pass, cw);
      side.route = pl->route;
      decode_entropy_pass<false>(ew, pkt, 256, m->st[s], pl, &side, &tape, data, &r, pass);
      decode_entropy_pass<true>(ew, pkt, 256, m->st[s], pl, &side, &tape, data, &r, pass);
      for (int c = 0; c < m->st[s]->channels; c++) decode_synth_pass(sw, m->st[s], pl, &side, epcm[s], c, pass);
      decode_plc_pass(pw, m->st[s], pl, &side, epcm[s], pass);
      decode_general_pass(gw, m->st[s], pl, &side, &tape, data, epcm[s], &r, pass, cw);
    }
  }
  if (ret > 0) {
    const int CH = m->L.nb_channels;
    for (int t = 0; t < ret; t++)
      for (int c = 0; c < CH; c++) {
        int mp = m->L.mapping[c];
        short v = 0;
        if (mp != 255) {
          if (mp < 2 * m->L.nb_coupled) v = epcm[mp >> 1][t * 2 + (mp & 1)];
          else v = epcm[mp - m->L.nb_coupled][t];
        }
        pcm[t * CH + c] = v;
      }
  }
  return ret;
}

}  // extern "C"
