// Rows A1 / N2: WAV ingest on the host (no CUDA in this file).
//
// The reference decodes one file per worker thread with torchaudio.load (tools/extract_embedding.py:26,
// tools/extract_speech_token.py:28); at 1,000 hours that is ~550 k files, and in Python the per-file interpreter work
// (open, parse, slice, copy) costs more than the GPU needs for the audio.  Two entry points take that off the
// interpreter: fy_wav_scan parses RIFF headers (durations for batch planning, the byte range of the samples) and
// fy_wav_read_s16 reads PCM16 payloads STRAIGHT INTO caller memory -- the pinned upload ring of pipeline.PinnedRing --
// with a pool of plain threads, one copy from the page cache to the buffer the DMA engine reads.  Files that are not
// plain PCM16 (other widths, float, extensible headers with odd layouts) are only flagged here and go through the
// Python reader (audio_io.py).
#include <errno.h>
#include <fcntl.h>
#include <string.h>
#include <sys/stat.h>
#include <unistd.h>

#include <atomic>
#include <thread>
#include <vector>

#include "fy_common.cuh"

namespace {

uint32_t rd32(const uint8_t* p) { return p[0] | (p[1] << 8) | (p[2] << 16) | ((uint32_t)p[3] << 24); }
uint16_t rd16(const uint8_t* p) { return (uint16_t)(p[0] | (p[1] << 8)); }

void scan_one(const char* path, fy_wav_info* o) {
  memset(o, 0, sizeof(*o));
  const int fd = open(path, O_RDONLY | O_CLOEXEC);
  if (fd < 0) { o->status = FY_WAV_ERR_OPEN; return; }
  uint8_t head[4096];
  const ssize_t got = pread(fd, head, sizeof(head), 0);
  struct stat st;
  const bool st_ok = fstat(fd, &st) == 0;
  close(fd);
  if (got < 12 || !st_ok) { o->status = FY_WAV_ERR_FORMAT; return; }
  if (memcmp(head, "RIFF", 4) != 0 || memcmp(head + 8, "WAVE", 4) != 0) { o->status = FY_WAV_ERR_FORMAT; return; }
  size_t pos = 12;
  bool have_fmt = false;
  while (pos + 8 <= (size_t)got) {
    const uint32_t csz = rd32(head + pos + 4);
    if (memcmp(head + pos, "fmt ", 4) == 0 && pos + 8 + 16 <= (size_t)got) {
      const uint8_t* f = head + pos + 8;
      o->format = rd16(f);
      o->channels = rd16(f + 2);
      o->sample_rate = (int32_t)rd32(f + 4);
      o->bits = rd16(f + 14);
      if (o->format == 0xFFFE && csz >= 26 && pos + 8 + 26 <= (size_t)got) o->format = rd16(f + 24);  // extensible: sub-format
      have_fmt = true;
    } else if (memcmp(head + pos, "data", 4) == 0) {
      if (!have_fmt || o->channels < 1 || o->bits < 8) break;
      const int64_t off = (int64_t)pos + 8;
      int64_t nbytes = csz;
      if (nbytes > (int64_t)st.st_size - off) nbytes = (int64_t)st.st_size - off;
      if (nbytes < 0) nbytes = 0;
      const int frame = o->channels * o->bits / 8;
      o->data_offset = off;
      o->frames = nbytes / (frame > 0 ? frame : 1);
      o->status = FY_WAV_OK;
      return;
    }
    pos += 8 + (size_t)csz + (csz & 1);
  }
  o->status = FY_WAV_ERR_LONG_HEADER;   // no data chunk inside the first 4 KiB (or fmt missing): let the Python reader decide
}

template <typename F>
void parallel_for(int n, int n_threads, F&& body) {
  if (n_threads < 1) n_threads = 1;
  if (n_threads > n) n_threads = n > 0 ? n : 1;
  std::atomic<int> next(0);
  auto work = [&]() {
    for (;;) {
      const int i = next.fetch_add(1, std::memory_order_relaxed);
      if (i >= n) return;
      body(i);
    }
  };
  std::vector<std::thread> pool;
  for (int t = 1; t < n_threads; ++t) pool.emplace_back(work);
  work();
  for (auto& t : pool) t.join();
}

}  // namespace

extern "C" {

int fy_wav_scan(const char* const* paths, int n, fy_wav_info* out, int n_threads) {
  FY_CHECK(paths && out && n >= 0, FY_ERR_ARG, "fy_wav_scan: bad argument");
  parallel_for(n, n_threads, [&](int i) { scan_one(paths[i], &out[i]); });
  return FY_OK;
}

int fy_wav_read_s16(const char* const* paths, const int64_t* data_offset, const int64_t* n_samples, int16_t* const* dst,
                    int n, int n_threads, int32_t* bad_index) {
  FY_CHECK(paths && data_offset && n_samples && dst && n >= 0, FY_ERR_ARG, "fy_wav_read_s16: bad argument");
  std::atomic<int> first_bad(-1);
  parallel_for(n, n_threads, [&](int i) {
    const int fd = open(paths[i], O_RDONLY | O_CLOEXEC);
    bool ok = fd >= 0;
    if (ok) {
      int64_t want = n_samples[i] * 2, done = 0;
      uint8_t* d = reinterpret_cast<uint8_t*>(dst[i]);
      while (done < want) {
        const ssize_t r = pread(fd, d + done, (size_t)(want - done), data_offset[i] + done);
        if (r <= 0) { if (r < 0 && errno == EINTR) continue; ok = false; break; }
        done += r;
      }
      close(fd);
    }
    if (!ok) { int exp = -1; first_bad.compare_exchange_strong(exp, i); }
  });
  const int bad = first_bad.load();
  if (bad_index) *bad_index = bad;
  FY_CHECK(bad < 0, FY_ERR_ARG, "fy_wav_read_s16: cannot read %lld samples of '%s'", (long long)n_samples[bad], paths[bad]);
  return FY_OK;
}

}  // extern "C"
