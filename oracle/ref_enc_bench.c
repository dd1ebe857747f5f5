/* ref_enc_bench.c -- threaded CPU baseline harness for the ENCODE path (TEST/BENCH INFRASTRUCTURE, ours).
 * Links the UNMODIFIED reference objects (oracle/_ref) and times opus_encode() on the host cores:
 * T threads, each owning n/T independent OpusEncoder objects (RESTRICTED_LOWDELAY = CELT-only),
 * fed from one raw int16 PCM file; stream i starts at frame (37*i) mod available, like the GPU batch.
 * Usage: ref_enc_bench <pcm.raw> <Fs> <channels> <bitrate> <complexity> <frame_size> <streams> <frames> <threads> [vbr]
 * (vbr: 0 = CBR, 1 = VBR, absent = the encoder's default).  Encoder objects are created before the clock starts.
 * Prints one line: stream_seconds wall_seconds threads checksum
 */
#include <pthread.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>
#include "opus.h"

static opus_int16 *g_pcm; static long g_avail_frames;
static int g_fs, g_ch, g_bitrate, g_cx, g_frame, g_frames, g_vbr = -1;
typedef struct { int first, last; uint32_t sum; double audio; OpusEncoder **enc; } Job;
static pthread_barrier_t g_start;

static void *worker(void *arg) {
  Job *job = (Job *)arg;
  int err, i, t, k;
  unsigned char pkt[1500];
  uint32_t sum = 0;
  job->enc = (OpusEncoder **)calloc(job->last - job->first + 1, sizeof(OpusEncoder *));
  for (i = job->first; i < job->last; i++) {
    OpusEncoder *enc = opus_encoder_create(g_fs, g_ch, OPUS_APPLICATION_RESTRICTED_LOWDELAY, &err);
    if (!enc) { fprintf(stderr, "create failed\n"); exit(1); }
    opus_encoder_ctl(enc, OPUS_SET_BITRATE(g_bitrate));
    opus_encoder_ctl(enc, OPUS_SET_COMPLEXITY(g_cx));
    if (g_vbr >= 0) opus_encoder_ctl(enc, OPUS_SET_VBR(g_vbr));
    job->enc[i - job->first] = enc;
  }
  pthread_barrier_wait(&g_start);
  pthread_barrier_wait(&g_start);
  for (i = job->first; i < job->last; i++) {
    OpusEncoder *enc = job->enc[i - job->first];
    for (t = 0; t < g_frames; t++) {
      const opus_int16 *in = g_pcm + (size_t)((37L * i + t) % g_avail_frames) * g_frame * g_ch;
      int ret = opus_encode(enc, in, g_frame, pkt, 1275);
      if (ret < 0) { fprintf(stderr, "encode failed %d\n", ret); exit(1); }
      for (k = 0; k < ret; k += 13) sum = sum * 31u + pkt[k];
      job->audio += (double)g_frame / g_fs;
    }
  }
  job->sum = sum;
  pthread_barrier_wait(&g_start);
  for (i = job->first; i < job->last; i++) opus_encoder_destroy(job->enc[i - job->first]);
  free(job->enc);
  return NULL;
}

int main(int argc, char **argv) {
  if (argc < 10) { fprintf(stderr, "usage: %s pcm.raw Fs ch bitrate complexity frame_size streams frames threads\n", argv[0]); return 2; }
  FILE *f = fopen(argv[1], "rb");
  if (!f) { perror(argv[1]); return 2; }
  g_fs = atoi(argv[2]); g_ch = atoi(argv[3]); g_bitrate = atoi(argv[4]); g_cx = atoi(argv[5]); g_frame = atoi(argv[6]);
  int streams = atoi(argv[7]); g_frames = atoi(argv[8]);
  int T = atoi(argv[9]);
  if (argc >= 11) g_vbr = atoi(argv[10]);
  fseek(f, 0, SEEK_END);
  long bytes = ftell(f);
  fseek(f, 0, SEEK_SET);
  g_pcm = (opus_int16 *)malloc(bytes);
  if (fread(g_pcm, 1, bytes, f) != (size_t)bytes) { fprintf(stderr, "short read\n"); return 2; }
  fclose(f);
  g_avail_frames = bytes / (2L * g_ch * g_frame);
  if (g_avail_frames < 1) { fprintf(stderr, "pcm file too small\n"); return 2; }
  if (T < 1) T = 1;
  if (T > streams) T = streams;
  pthread_t *th = (pthread_t *)malloc(sizeof(pthread_t) * T);
  Job *jobs = (Job *)calloc(T, sizeof(Job));
  struct timespec t0, t1;
  pthread_barrier_init(&g_start, NULL, T + 1);
  for (int i = 0; i < T; i++) {
    jobs[i].first = (int)((long long)streams * i / T);
    jobs[i].last = (int)((long long)streams * (i + 1) / T);
    pthread_create(&th[i], NULL, worker, &jobs[i]);
  }
  pthread_barrier_wait(&g_start);
  clock_gettime(CLOCK_MONOTONIC, &t0);
  pthread_barrier_wait(&g_start);
  pthread_barrier_wait(&g_start);
  clock_gettime(CLOCK_MONOTONIC, &t1);
  double audio = 0; uint32_t sum = 0;
  for (int i = 0; i < T; i++) { pthread_join(th[i], NULL); audio += jobs[i].audio; sum ^= jobs[i].sum; }
  double wall = (t1.tv_sec - t0.tv_sec) + 1e-9 * (t1.tv_nsec - t0.tv_nsec);
  printf("%.6f %.6f %d %08x\n", audio, wall, T, sum);
  return 0;
}
