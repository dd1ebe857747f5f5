// tests/emu/emu_enc.cpp -- TEST-ONLY build of the encoder kernel bodies as plain C++ (CB_HOST_EMU),
// the encode-side twin of emu_celt.cpp.  Lets the CPU suite compare packets with the compiled
// reference without a GPU.  Never linked into libopus_b200.so.
#define CB_HOST_EMU 1
#include <stdlib.h>
#include <string.h>
#include "../../concentus_b200/csrc/cb_opus_enc.h"
#include "../../concentus_b200/csrc/cb_enc_ctl.h"
#include "../../concentus_b200/csrc/cb_surround.h"

extern "C" {

void* emu_enc_create(int Fs, int channels, int application) {
  if ((Fs != 48000 && Fs != 24000 && Fs != 16000 && Fs != 12000 && Fs != 8000) || (channels != 1 && channels != 2) ||
      (application != CB_APP_VOIP && application != CB_APP_AUDIO && application != CB_APP_LOWDELAY))
    return nullptr;
  OpusEncState* s = (OpusEncState*)malloc(sizeof(OpusEncState));
  opus_enc_init(s, Fs, channels, application);
  return s;
}
void emu_enc_destroy(void* s) { free(s); }
unsigned emu_enc_final_range(void* s) { return ((OpusEncState*)s)->o.rangeFinal; }

int emu_enc_ctl(void* state, int request, int value, int* out) {
  OpusEncState* st = (OpusEncState*)state;
  if (request == CB_REQ_RESET_STATE) { opus_enc_reset(st); return CB_OK; }
  i32 tmp = 0;
  int rc = opus_enc_ctl_scalars(&st->o, &st->celt.s, request, value, out ? (i32*)out : &tmp);
  return rc;
}

// OPUS_SET_ENERGY_MASK: 21 values per channel, or NULL
void emu_enc_set_mask(void* state, const short* mask) {
  OpusEncState* st = (OpusEncState*)state;
  st->celt.s.has_mask = mask ? 1 : 0;
  if (mask) memcpy(st->celt.energy_mask, mask, sizeof(i16) * CB_NBANDS * st->o.channels);
}

int emu_enc_encode(void* state, const short* pcm, int frame_size, unsigned char* data, int max_bytes) {
  OpusEncState* st = (OpusEncState*)state;
  static EncWork w;
  static EncScratch sc;
  return opus_encode_stream<true>(w, sc, st, pcm, frame_size, data, max_bytes);
}

// surround_analysis: pcm interleaved by `channels`; mem = channels*120 window memory, preemph_mem = channels values
// (both updated); out = 21 masks per channel
void emu_surround_analysis(const short* pcm, int channels, int len, int Fs, int* mem, int* preemph_mem, short* out) {
  static SurroundState st;
  static SurroundScratch sc;
  for (int c = 0; c < channels; c++) {
    memcpy(st.window_mem[c], mem + c * CB_OVERLAP, sizeof(i32) * CB_OVERLAP);
    st.preemph_mem[c] = preemph_mem[c];
  }
  for (int c = 0; c < channels; c++) surround_channel(pcm, channels, c, len, Fs, &st, sc, out + CB_NBANDS * c);
  surround_combine(out, channels);
  for (int c = 0; c < channels; c++) {
    memcpy(mem + c * CB_OVERLAP, st.window_mem[c], sizeof(i32) * CB_OVERLAP);
    preemph_mem[c] = st.preemph_mem[c];
  }
}

}  // extern "C"
