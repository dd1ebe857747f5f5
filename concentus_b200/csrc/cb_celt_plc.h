// cb_celt_plc.h -- packet-loss concealment for CELT frames (kernel C body + the noise branch used
// by kernel A).  Bit-exact with celt_decode_lost() (opus-fix/celt/celt_decoder.c:415-711),
// celt_plc_pitch_search (:399-413), pitch_downsample / pitch_search / find_best_pitch / celt_fir5
// (celt/pitch.c:45-370), _celt_autocorr / _celt_lpc / celt_fir / celt_iir (celt/celt_lpc.c:37-328).
//
// One warp per stream.  Concealment is rare (it runs only for lost or DTX frames), so the layout
// favours clarity: channels are processed one after the other, inner products are lane loops folded
// with redux, and the two truly serial recurrences (Levinson-Durbin, the order-24 IIR) keep their
// order -- the IIR spreads its 24 taps over lanes and folds once per sample.
#pragma once
#include "cb_common.h"
#include "cb_state.h"
#include "cb_celt_synth.h"
#include "cb_celt_entropy.h"
#include "cb_pitch.h"

#define CB_MAX_PERIOD 1024
#define CB_PLC_PITCH_LAG_MAX 720
#define CB_PLC_PITCH_LAG_MIN 100

struct alignas(16) PlcWork {
  i32 nb[CB_MAXN + CB_OVERLAP];      // D_new[2048-N .. 2048+overlap): the frame being concealed
  i32 etmp[CB_OVERLAP];
  i32 xcorr[(CB_PLC_PITCH_LAG_MAX - CB_PLC_PITCH_LAG_MIN) >> 1];
  i32 ac[CB_LPC_ORDER + 1];
  i32 lpc32[CB_LPC_ORDER];
  i16 exc[CB_MAX_PERIOD + 32];       // excitation; [CB_MAX_PERIOD..) is IIR history scratch
  i16 xx[CB_MAX_PERIOD];             // windowed / scaled copy (autocorr), FIR output staging
  i16 xlp[CB_HIST >> 1];             // 2x decimated pitch buffer
  i16 xlp4[(CB_HIST - CB_PLC_PITCH_LAG_MAX) >> 2];
  i16 ylp4[CB_HIST >> 2];
  i16 lpcmem[CB_LPC_ORDER];
  i16 yhist[CB_MAXN + CB_OVERLAP + CB_LPC_ORDER];   // rounded IIR outputs (filter memory)
};

CB_FN i32 hist_at(const i32* hist, int pos, int i) { return hist[(pos + i) & (CB_HIST - 1)]; }

// celt_plc_pitch_search (celt_decoder.c:399-413) = pitch_downsample + pitch_search
CB_FN int plc_pitch_search(PlcWork& w, const CeltDecState* st) {
  const int C = st->channels;
  const int len = CB_HIST;
  i16* x_lp = w.xlp;
  // ---- pitch_downsample (pitch.c:147-222) ----
  i32 mx = 0, mn = 0;
  CB_NOUNROLL for (int c = 0; c < C; c++) {
    const i32* h = st->hist[c];
    const int pos = st->hist_pos[c];
    CB_LANE_LOOP(i, len) {
      i32 v = hist_at(h, pos, i);
      mx = max32(mx, v);
      mn = min32(mn, v);
    }
  }
  mx = cb_wmax(mx);
  mn = cb_wmin(mn);
  i32 maxabs = max32(mx, neg32(mn));
  if (maxabs < 1) maxabs = 1;
  int shift = cb_ilog2(maxabs) - 10;
  if (shift < 0) shift = 0;
  if (C == 2) shift++;
  CB_WSYNC();
  CB_LANE_LOOP(i, len >> 1) {
    i32 acc = 0;
    for (int c = 0; c < C; c++) {
      const i32* h = st->hist[c];
      const int pos = st->hist_pos[c];
      i32 v;
      if (i == 0) v = half32(add32(half32(hist_at(h, pos, 1)), hist_at(h, pos, 0))) >> shift;
      else v = half32(add32(half32(add32(hist_at(h, pos, 2 * i - 1), hist_at(h, pos, 2 * i + 1))), hist_at(h, pos, 2 * i))) >> shift;
      // x_lp is opus_val16: the first channel is stored (truncated), the second added in 16 bits
      acc = (c == 0) ? (i32)(i16)v : (i32)(i16)(acc + v);
    }
    x_lp[i] = (i16)acc;
  }
  CB_WSYNC();
  pitch_whiten(x_lp, len >> 1, w.ac, w.lpc32, w.xx);
  int pitch = pitch_search(x_lp + (CB_PLC_PITCH_LAG_MAX >> 1), x_lp, CB_HIST - CB_PLC_PITCH_LAG_MAX,
                           CB_PLC_PITCH_LAG_MAX - CB_PLC_PITCH_LAG_MIN, w.xlp4, w.ylp4, w.xcorr);
  return CB_PLC_PITCH_LAG_MAX - pitch;
}

// Pitch-based concealment of one frame of N samples (celt_decoder.c:507-706) + de-emphasis.
CB_FN void celt_plc_pitch(PlcWork& w, CeltDecState* st, int N, i16* pcm, int accum, int decode_gain = 0) {
  const int C = st->channels;
  const int loss_count = st->loss_count;
  i32 fade = CB_Q15ONE;
  int pitch_index;
  if (loss_count == 0) {
    pitch_index = plc_pitch_search(w, st);
    st->last_pitch_index = pitch_index;
  } else {
    pitch_index = st->last_pitch_index;
    fade = 26214;                                      // QCONST16(.8f,15)
  }
  CB_NOUNROLL for (int c = 0; c < C; c++) {
    const i32* hist = st->hist[c];
    const int pos = st->hist_pos[c];                   // D_old[i] = hist[(pos+i)&2047]
    i16* exc = w.exc;
    i16* lpc = st->lpc + c * CB_LPC_ORDER;
    CB_WSYNC();
    CB_LANE_LOOP(i, CB_MAX_PERIOD) exc[i] = round16(hist_at(hist, pos, CB_HIST - CB_MAX_PERIOD + i), CB_SIG_SHIFT);
    CB_WSYNC();
    if (loss_count == 0) {
      pitch_autocorr(exc, w.ac, w.xx, 1, CB_LPC_ORDER, CB_MAX_PERIOD);
      i32 a0 = add32(w.ac[0], w.ac[0] >> 13);
      CB_WSYNC();
      CB_LANE_LOOP(i, CB_LPC_ORDER + 1) {
        if (i == 0) w.ac[0] = a0;
        else w.ac[i] = sub32(w.ac[i], mult16_32_q15(2 * i * i, w.ac[i]));
      }
      CB_WSYNC();
      i16 l[CB_LPC_ORDER];
      pitch_lpc(l, w.ac, w.lpc32, CB_LPC_ORDER);
      CB_NOUNROLL for (int i = 0; i < CB_LPC_ORDER; i++) lpc[i] = l[i];
    }
    CB_WSYNC();
    const int exc_length = cb_imin(2 * pitch_index, CB_MAX_PERIOD);
    // celt_fir (celt_lpc.c:95-149) on the last exc_length samples; memory = samples just before
    CB_LANE_LOOP(i, CB_LPC_ORDER) w.lpcmem[i] = round16(hist_at(hist, pos, CB_HIST - exc_length - 1 - i), CB_SIG_SHIFT);
    CB_WSYNC();
    {
      i16* x = exc + CB_MAX_PERIOD - exc_length;
      CB_LANE_LOOP(i, exc_length) {
        i32 sum = 0;
        for (int j = 0; j < CB_LPC_ORDER; j++) {
          int k = i - 1 - j;
          i32 xv = k >= 0 ? x[k] : w.lpcmem[-k - 1];
          sum = mac16_16(sum, lpc[j], xv);
        }
        w.xx[i] = sat16(add32((i32)x[i], pshr32(sum, CB_SIG_SHIFT)));
      }
      CB_WSYNC();
      CB_LANE_LOOP(i, exc_length) x[i] = w.xx[i];
      CB_WSYNC();
    }
    // decay of the excitation energy over the last two pitch periods
    i32 decay;
    {
      i32 mx = 0, mn = 0;
      CB_LANE_LOOP(i, exc_length) { i32 v = exc[CB_MAX_PERIOD - exc_length + i]; mx = max32(mx, v); mn = min32(mn, v); }
      i32 maxabs = max32(cb_wmax(mx), neg32(cb_wmin(mn)));
      int shift = cb_imax(0, 2 * cb_zlog2(maxabs) - 20);
      int decay_length = exc_length >> 1;
      i32 e1 = 0, e2 = 0;
      CB_LANE_LOOP(i, decay_length) {
        i32 e = exc[CB_MAX_PERIOD - decay_length + i];
        e1 = add32(e1, mult16_16(e, e) >> shift);
        e = exc[CB_MAX_PERIOD - 2 * decay_length + i];
        e2 = add32(e2, mult16_16(e, e) >> shift);
      }
      i32 E1 = add32(1, cb_wsum(e1)), E2 = add32(1, cb_wsum(e2));
      E1 = min32(E1, E2);
      decay = (i16)cb_sqrt(cb_frac_div32(E1 >> 1, E2));
    }
    // extrapolate the excitation with period pitch_index, each period attenuated by `decay`
    const int extrapolation_offset = CB_MAX_PERIOD - pitch_index;
    const int extrapolation_len = N + CB_OVERLAP;
    i32 S1p = 0;
    CB_WSYNC();
    CB_LANE_LOOP(i, extrapolation_len) {
      int k = i / pitch_index, j = i - k * pitch_index;
      i32 att = (i16)mult16_16_q15(fade, decay);
      for (int t = 0; t < k; t++) att = (i16)mult16_16_q15(att, decay);
      w.nb[i] = shl32((i32)(i16)mult16_16_q15(att, exc[extrapolation_offset + j]), CB_SIG_SHIFT);
      i32 tmp = round16(hist_at(hist, pos, CB_HIST - CB_MAX_PERIOD + extrapolation_offset + j), CB_SIG_SHIFT);
      S1p = add32(S1p, mult16_16(tmp, tmp) >> 8);
    }
    i32 S1 = cb_wsum(S1p);
    CB_WSYNC();
    // celt_iir (celt_lpc.c:151-230): y[i] = x[i] - sum_j lpc[j]*round16(y[i-1-j]); memory = last samples
    {
      i16* yh = w.yhist;                               // yh[CB_LPC_ORDER + i] = round16(y[i])
      CB_LANE_LOOP(i, CB_LPC_ORDER) yh[CB_LPC_ORDER - 1 - i] = round16(hist_at(hist, pos, CB_HIST - 1 - i), CB_SIG_SHIFT);
      CB_WSYNC();
      CB_NOUNROLL for (int i = 0; i < extrapolation_len; i++) {
        i32 part = 0;
        CB_LANE_LOOP(j, CB_LPC_ORDER) part = sub32(part, mult16_16(lpc[j], yh[CB_LPC_ORDER + i - 1 - j]));
        i32 sum = add32(w.nb[i], cb_wsum(part));
        CB_WSYNC();
        w.nb[i] = sum;
        yh[CB_LPC_ORDER + i] = round16(sum, CB_SIG_SHIFT);
        CB_WSYNC();
      }
    }
    // energy guard
    {
      i32 s2 = 0;
      CB_LANE_LOOP(i, extrapolation_len) {
        i32 tmp = round16(w.nb[i], CB_SIG_SHIFT);
        s2 = add32(s2, mult16_16(tmp, tmp) >> 8);
      }
      i32 S2 = cb_wsum(s2);
      CB_WSYNC();
      if (!(S1 > (S2 >> 2))) {
        CB_LANE_LOOP(i, extrapolation_len) w.nb[i] = 0;
      } else if (S1 < S2) {
        i32 ratio = (i16)cb_sqrt(cb_frac_div32(add32(S1 >> 1, 1), add32(S2, 1)));
        CB_LANE_LOOP(i, extrapolation_len) {
          if (i < CB_OVERLAP) {
            i32 tmp_g = (i16)(CB_Q15ONE - mult16_16_q15(cb_ldg(&ct_window120[i]), CB_Q15ONE - ratio));
            w.nb[i] = mult16_32_q15(tmp_g, w.nb[i]);
          } else {
            w.nb[i] = mult16_32_q15(ratio, w.nb[i]);
          }
        }
      }
      CB_WSYNC();
    }
    // pre-filter the overlap for the next frame (comb_filter with x != y, negated gain) + TDAC
    {
      const int T = st->pf_period;
      const i32 g = (i16)(-st->pf_gain);
      const int tap = st->pf_tapset;
      if (g == 0) {
        CB_LANE_LOOP(i, CB_OVERLAP) w.etmp[i] = w.nb[N + i];
      } else {
        i32 g10 = mult16_16_p15(g, ct_comb_gains[tap][0]);
        i32 g11 = mult16_16_p15(g, ct_comb_gains[tap][1]);
        i32 g12 = mult16_16_p15(g, ct_comb_gains[tap][2]);
        // x[k] = D_new[2048+k] = nb[N+k] (k >= -N) else the ring (D_new[m] = D_old[m+N])
        CB_LANE_LOOP(i, CB_OVERLAP) {
          i32 v[5];
          for (int d = 0; d < 5; d++) {
            int k = i - T + 2 - d;                     // x[i-T+2], x[i-T+1], x[i-T], x[i-T-1], x[i-T-2]
            v[d] = (k >= -N) ? w.nb[N + k] : hist_at(hist, pos, CB_HIST + N + k);
          }
          i32 yv = w.nb[N + i];
          yv = add32(yv, mult16_32_q15(g10, v[2]));
          yv = add32(yv, mult16_32_q15(g11, add32(v[1], v[3])));
          yv = add32(yv, mult16_32_q15(g12, add32(v[0], v[4])));
          w.etmp[i] = yv;
        }
      }
      CB_WSYNC();
      CB_LANE_LOOP(i, CB_OVERLAP / 2) {
        i32 w0 = cb_ldg(&ct_window120[i]), w1 = cb_ldg(&ct_window120[CB_OVERLAP - i - 1]);
        st->tail[c][i] = add32(mult16_32_q15(w0, w.etmp[CB_OVERLAP - 1 - i]), mult16_32_q15(w1, w.etmp[i]));
      }
      CB_WSYNC();
    }
    emit_frame(w.nb, N, st, c, pcm, accum, decode_gain);
  }
  CB_WSYNC();
  st->loss_count = loss_count + 1;
}

// (noise-based concealment, the other branch of celt_decode_lost: celt_plc_noise_parse / _rebuild in cb_celt_entropy.h)
