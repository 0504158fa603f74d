// BGZF (blocked gzip) reader and writer over zlib.  Host-side I/O of the drop-in CLI; htslib/libdeflate are
// not in this image, so the codec is written here (same container format: 64 KiB blocks, BC extra field,
// EOF marker block).  Reader and writer (de)compress the independent blocks on small thread pools, keeping block order.
#pragma once
#include <zlib.h>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <condition_variable>
#include <deque>
#include <functional>
#include <mutex>
#include <stdexcept>
#include <string>
#include <thread>
#include <vector>

namespace brbhost {

// Reader.  threads <= 1: blocks are inflated on the calling thread.  threads > 1: a feeder thread walks the file
// block by block (the BC field gives every block's size without touching the payload), a pool inflates the
// independent blocks, read() consumes them in file order.
// With a piece inflater (the engine's brb_bgzf_inflate): the feeder cuts the file into PIECES of whole blocks (1 MiB first, for a quick
// start, then 8 MiB), two workers hand them to the inflater -- which may decline (returns false): the piece is then inflated here,
// block by block, and that is also where a corrupt block gets its error message.
class BgzfReader {
 public:
  using PieceInflater = std::function<bool(int worker, const uint8_t* bgzf, size_t n, std::vector<uint8_t>& out)>;
  explicit BgzfReader(const std::string& path, int threads = 1, PieceInflater ext = nullptr) : ext_(std::move(ext)) {
    f_ = fopen(path.c_str(), "rb");
    if (!f_) throw std::runtime_error("cannot open " + path);
    setvbuf(f_, nullptr, _IOFBF, 1 << 20);
    if (ext_) {
      max_jobs_ = 4;
      for (int i = 0; i < 2; i++) pool_.emplace_back([this, i] { worker(i); });
      feeder_ = std::thread([this] { feed_pieces(); });
    } else if (threads > 1) {
      max_jobs_ = (size_t)threads * 8;
      for (int i = 0; i < threads; i++) pool_.emplace_back([this] { worker(0); });
      feeder_ = std::thread([this] { feed(); });
    }
  }
  ~BgzfReader() {
    { std::unique_lock<std::mutex> lk(m_); stop_ = true; }
    cv_.notify_all(); cv_room_.notify_all(); cv_ready_.notify_all();
    if (feeder_.joinable()) feeder_.join();
    for (auto& t : pool_) t.join();
    for (Job* j : order_) delete j;
    for (Job* j : free_) delete j;
    if (f_) fclose(f_);
  }
  // reads exactly n bytes unless EOF; returns bytes read
  size_t read(void* dst, size_t n) {
    uint8_t* d = static_cast<uint8_t*>(dst); size_t got = 0;
    while (got < n) {
      if (pos_ == buf_.size()) { if (!(pool_.empty() ? fill() : next_block())) break; }
      size_t k = std::min(n - got, buf_.size() - pos_);
      memcpy(d + got, buf_.data() + pos_, k); pos_ += k; got += k;
    }
    return got;
  }
  // The rest of the current inflated block, or the whole next one (false at EOF): the record walk of the CLI takes the stream block by
  // block instead of paying two read() calls per record.  The bytes stay valid until the next call.
  bool next_chunk(const uint8_t*& p, size_t& n) {
    if (pos_ == buf_.size()) { if (!(pool_.empty() ? fill() : next_block())) return false; }
    p = buf_.data() + pos_; n = buf_.size() - pos_; pos_ = buf_.size();
    return true;
  }
 private:
  struct Job { std::vector<uint8_t> comp, out; size_t clen = 0; uint32_t isize = 0; bool ready = false, piece = false; std::string err; };
  // next compressed block from the file: false at EOF
  bool read_block(std::vector<uint8_t>& comp, size_t& clen, uint32_t& isize) {
    uint8_t h[12];
    if (fread(h, 1, 12, f_) != 12) return false;
    if (h[0] != 0x1f || h[1] != 0x8b) throw std::runtime_error("not a BGZF file");
    const unsigned xlen = h[10] | (h[11] << 8);
    extra_.resize(xlen);
    if (xlen && fread(extra_.data(), 1, xlen, f_) != xlen) return false;
    unsigned bsize = 0; bool found = false;
    for (size_t p = 0; p + 4 <= xlen;) {
      const unsigned slen = extra_[p + 2] | (extra_[p + 3] << 8);
      if (extra_[p] == 'B' && extra_[p + 1] == 'C' && slen == 2) { bsize = extra_[p + 4] | (extra_[p + 5] << 8); found = true; }
      p += 4 + slen;
    }
    if (!found) throw std::runtime_error("BGZF block without BC field");
    clen = (size_t)bsize + 1 - xlen - 12 - 8;
    comp.resize(clen + 8);
    if (fread(comp.data(), 1, clen + 8, f_) != clen + 8) return false;
    isize = comp[clen + 4] | (comp[clen + 5] << 8) | (comp[clen + 6] << 16) | ((uint32_t)comp[clen + 7] << 24);
    return true;
  }
  static void inflate_block(const std::vector<uint8_t>& comp, size_t clen, uint32_t isize, std::vector<uint8_t>& out) {
    out.resize(isize);
    if (isize == 0) return;
    z_stream zs; memset(&zs, 0, sizeof zs);
    if (inflateInit2(&zs, -15) != Z_OK) throw std::runtime_error("inflateInit2");
    zs.next_in = const_cast<uint8_t*>(comp.data()); zs.avail_in = (uInt)clen; zs.next_out = out.data(); zs.avail_out = isize;
    const int rc = inflate(&zs, Z_FINISH); inflateEnd(&zs);
    if (rc != Z_STREAM_END || zs.total_out != isize) throw std::runtime_error("BGZF inflate failed");
    const uint32_t want = comp[clen] | (comp[clen + 1] << 8) | (comp[clen + 2] << 16) | ((uint32_t)comp[clen + 3] << 24);
    if ((uint32_t)crc32(crc32(0, nullptr, 0), out.data(), isize) != want) throw std::runtime_error("BGZF block CRC mismatch (corrupt input)");
  }
  bool fill() {
    for (;;) {
      size_t clen; uint32_t isize;
      if (!read_block(comp_, clen, isize)) return false;
      pos_ = 0;
      inflate_block(comp_, clen, isize, buf_);
      if (isize) return true;
    }
  }
  // ---- threaded path
  void feed() {
    try {
      for (;;) {
        Job* j = take_job();
        if (!read_block(j->comp, j->clen, j->isize)) { delete j; break; }
        std::unique_lock<std::mutex> lk(m_);
        cv_room_.wait(lk, [this] { return stop_ || order_.size() < max_jobs_; });
        if (stop_) { delete j; return; }
        order_.push_back(j); todo_.push_back(j);
        cv_.notify_one();
      }
    } catch (const std::exception& e) { std::unique_lock<std::mutex> lk(m_); feed_err_ = e.what(); }
    { std::unique_lock<std::mutex> lk(m_); eof_ = true; }
    cv_.notify_all(); cv_ready_.notify_all();
  }
  // one whole block (header, payload, footer) appended to dst; false at EOF
  bool read_raw_block(std::vector<uint8_t>& dst) {
    uint8_t h[12];
    if (fread(h, 1, 12, f_) != 12) return false;
    if (h[0] != 0x1f || h[1] != 0x8b) throw std::runtime_error("not a BGZF file");
    const unsigned xlen = h[10] | (h[11] << 8);
    const size_t at = dst.size();
    dst.resize(at + 12 + xlen);
    memcpy(dst.data() + at, h, 12);
    if (xlen && fread(dst.data() + at + 12, 1, xlen, f_) != xlen) { dst.resize(at); return false; }
    unsigned bsize = 0; bool found = false;
    for (size_t p = 0; p + 4 <= xlen;) {
      const uint8_t* x = dst.data() + at + 12 + p; const unsigned slen = x[2] | (x[3] << 8);
      if (x[0] == 'B' && x[1] == 'C' && slen == 2) { bsize = x[4] | (x[5] << 8); found = true; }
      p += 4 + slen;
    }
    if (!found || (size_t)bsize + 1 < 12 + (size_t)xlen + 8) throw std::runtime_error("BGZF block without BC field");
    const size_t rest = (size_t)bsize + 1 - 12 - xlen;
    dst.resize(at + 12 + xlen + rest);
    if (fread(dst.data() + at + 12 + xlen, 1, rest, f_) != rest) { dst.resize(at); return false; }
    return true;
  }
  // Block jobs are recycled: a fresh pair of 64 KiB vectors per block means page faults and allocator traffic on every pool thread.
  Job* take_job() {
    { std::unique_lock<std::mutex> lk(m_); if (!free_.empty()) { Job* j = free_.back(); free_.pop_back(); j->ready = false; j->err.clear(); return j; } }
    return new Job();
  }
  void give_job(Job* j) { std::unique_lock<std::mutex> lk(m_); if (free_.size() < max_jobs_ + 8) free_.push_back(j); else { lk.unlock(); delete j; } }
  void feed_pieces() {
    try {
      size_t want = 1 << 20; bool more = true;
      while (more) {
        Job* j = new Job();
        j->comp.reserve(want + (64 << 10));
        while (j->comp.size() < want && (more = read_raw_block(j->comp))) {}
        j->clen = j->comp.size(); j->piece = true;
        if (j->clen == 0) { delete j; break; }
        want = 8 << 20;
        std::unique_lock<std::mutex> lk(m_);
        cv_room_.wait(lk, [this] { return stop_ || order_.size() < max_jobs_; });
        if (stop_) { delete j; return; }
        order_.push_back(j); todo_.push_back(j);
        cv_.notify_one();
      }
    } catch (const std::exception& e) { std::unique_lock<std::mutex> lk(m_); feed_err_ = e.what(); }
    { std::unique_lock<std::mutex> lk(m_); eof_ = true; }
    cv_.notify_all(); cv_ready_.notify_all();
  }
  // a piece on the host: its blocks one after the other
  static void inflate_piece(const std::vector<uint8_t>& piece, std::vector<uint8_t>& out) {
    out.clear();
    std::vector<uint8_t> comp, one;
    for (size_t p = 0; p < piece.size();) {
      const unsigned xlen = piece[p + 10] | (piece[p + 11] << 8);
      unsigned bsize = 0;
      for (size_t q = 0; q + 4 <= xlen;) { const uint8_t* x = piece.data() + p + 12 + q; const unsigned slen = x[2] | (x[3] << 8); if (x[0] == 'B' && x[1] == 'C' && slen == 2) bsize = x[4] | (x[5] << 8); q += 4 + slen; }
      const size_t end = p + bsize + 1, clen = end - 8 - (p + 12 + xlen);
      comp.assign(piece.begin() + p + 12 + xlen, piece.begin() + end);
      const uint32_t isize = comp[clen + 4] | (comp[clen + 5] << 8) | (comp[clen + 6] << 16) | ((uint32_t)comp[clen + 7] << 24);
      inflate_block(comp, clen, isize, one);
      out.insert(out.end(), one.begin(), one.end());
      p = end;
    }
  }
  void worker(int id) {
    for (;;) {
      Job* j = nullptr;
      { std::unique_lock<std::mutex> lk(m_); cv_.wait(lk, [this] { return stop_ || eof_ || !todo_.empty(); });
        if (todo_.empty()) { if (stop_ || eof_) return; continue; }
        j = todo_.front(); todo_.pop_front(); }
      try {
        if (!j->piece) inflate_block(j->comp, j->clen, j->isize, j->out);
        else if (!ext_(id, j->comp.data(), j->clen, j->out)) inflate_piece(j->comp, j->out);
      } catch (const std::exception& e) { j->err = e.what(); }
      { std::unique_lock<std::mutex> lk(m_); j->ready = true; }
      cv_ready_.notify_all();
    }
  }
  bool next_block() {
    for (;;) {
      Job* j = nullptr;
      { std::unique_lock<std::mutex> lk(m_);
        cv_ready_.wait(lk, [this] { return (!order_.empty() && order_.front()->ready) || (order_.empty() && eof_); });
        if (order_.empty()) { if (!feed_err_.empty()) throw std::runtime_error(feed_err_); return false; }
        j = order_.front(); order_.pop_front(); }
      cv_room_.notify_one();
      if (!j->err.empty()) { const std::string e = j->err; delete j; throw std::runtime_error(e); }
      buf_.swap(j->out); pos_ = 0;
      if (j->piece) delete j; else give_job(j);             // (block jobs go round: their two 64 KiB buffers keep their pages)
      if (!buf_.empty()) return true;
    }
  }
  FILE* f_ = nullptr; std::vector<uint8_t> buf_, comp_, extra_; size_t pos_ = 0;
  std::vector<std::thread> pool_; std::thread feeder_; std::mutex m_; std::condition_variable cv_, cv_room_, cv_ready_;
  std::deque<Job*> order_, todo_; std::vector<Job*> free_; size_t max_jobs_ = 0; bool stop_ = false, eof_ = false; std::string feed_err_;
  PieceInflater ext_;
};

class BgzfWriter {
 public:
  static constexpr size_t BLOCK = 0xff00;
  BgzfWriter(const std::string& path, int level, int threads) : level_(level) {
    f_ = fopen(path.c_str(), "wb");
    if (!f_) throw std::runtime_error("cannot create " + path);
    setvbuf(f_, nullptr, _IOFBF, 1 << 20);
    if (threads < 1) threads = 1;
    for (int i = 0; i < threads; i++) pool_.emplace_back([this] { worker(); });
  }
  ~BgzfWriter() { try { close(); } catch (...) {} }
  void write(const void* src, size_t n) {
    const uint8_t* s = static_cast<const uint8_t*>(src);
    while (n) {
      size_t k = std::min(n, BLOCK - cur_.size());
      cur_.insert(cur_.end(), s, s + k); s += k; n -= k;
      if (cur_.size() == BLOCK) submit();
    }
  }
  void flush_block() { if (!cur_.empty()) submit(); }   // htslib flushes after the header
  // ready-made BGZF blocks (compressed elsewhere: the engine's device encoder), appended after everything written so far
  void write_raw(const void* src, size_t n) {
    flush_block(); drain(true);
    if (n && fwrite(src, 1, n, f_) != n) fail("short write");
  }
  void close() {
    if (!f_) return;
    flush_block();
    { std::unique_lock<std::mutex> lk(m_); done_ = true; }
    cv_.notify_all();
    for (auto& t : pool_) t.join();
    pool_.clear();
    drain(true);
    static const uint8_t eof[28] = {0x1f, 0x8b, 8, 4, 0, 0, 0, 0, 0, 0xff, 6, 0, 'B', 'C', 2, 0, 0x1b, 0, 3, 0, 0, 0, 0, 0, 0, 0, 0, 0};
    const bool ok = fwrite(eof, 1, 28, f_) == 28 && !failed_;
    FILE* f = f_; f_ = nullptr;
    if (fclose(f) != 0 || !ok) throw std::runtime_error("writing the output BAM failed (" + (err_.empty() ? std::string("I/O error, disk full?") : err_) + ")");
  }
 private:
  struct Job { std::vector<uint8_t> in, out; bool ready = false; };
  void submit() {
    Job* j = new Job(); j->in.swap(cur_); cur_.reserve(BLOCK);
    { std::unique_lock<std::mutex> lk(m_); order_.push_back(j); todo_.push_back(j); }
    cv_.notify_one();
    drain(false);
    // bound memory: wait when too many blocks are in flight
    std::unique_lock<std::mutex> lk(m_);
    while (order_.size() > 256) { lk.unlock(); drain(false); std::this_thread::yield(); lk.lock(); }
  }
  void worker() {
    for (;;) {
      Job* j = nullptr;
      { std::unique_lock<std::mutex> lk(m_); cv_.wait(lk, [this] { return done_ || !todo_.empty(); }); if (todo_.empty()) return; j = todo_.front(); todo_.pop_front(); }
      compress(*j);
      { std::unique_lock<std::mutex> lk(m_); j->ready = true; }
    }
  }
  void compress(Job& j) {
    j.out.resize(compressBound((uLong)j.in.size()) + 32);
    z_stream zs; memset(&zs, 0, sizeof zs);
    if (deflateInit2(&zs, level_, Z_DEFLATED, -15, 8, Z_DEFAULT_STRATEGY) != Z_OK) { fail("deflateInit2 failed"); j.out.clear(); return; }
    zs.next_in = j.in.data(); zs.avail_in = (uInt)j.in.size(); zs.next_out = j.out.data() + 18; zs.avail_out = (uInt)(j.out.size() - 26);
    const int rc = deflate(&zs, Z_FINISH); const size_t clen = zs.total_out; deflateEnd(&zs);
    if (rc != Z_STREAM_END) { fail("deflate failed"); j.out.clear(); return; }
    static const uint8_t head[16] = {0x1f, 0x8b, 8, 4, 0, 0, 0, 0, 0, 0xff, 6, 0, 'B', 'C', 2, 0};
    memcpy(j.out.data(), head, 16);
    const uint16_t bs = (uint16_t)(clen + 25); memcpy(j.out.data() + 16, &bs, 2);
    const uint32_t crc = (uint32_t)crc32(crc32(0, nullptr, 0), j.in.data(), (uInt)j.in.size()), isz = (uint32_t)j.in.size();
    memcpy(j.out.data() + 18 + clen, &crc, 4); memcpy(j.out.data() + 22 + clen, &isz, 4);
    j.out.resize(clen + 26);
  }
  void drain(bool all) {
    for (;;) {
      Job* j = nullptr;
      { std::unique_lock<std::mutex> lk(m_); if (order_.empty()) return; if (!order_.front()->ready) { if (!all) return; lk.unlock(); std::this_thread::yield(); continue; } j = order_.front(); order_.pop_front(); }
      if (fwrite(j->out.data(), 1, j->out.size(), f_) != j->out.size()) fail("short write");
      delete j;
    }
  }
  void fail(const char* what) { std::unique_lock<std::mutex> lk(m_); failed_ = true; if (err_.empty()) err_ = what; }
  bool failed_ = false; std::string err_;
  FILE* f_ = nullptr; int level_; std::vector<uint8_t> cur_;
  std::vector<std::thread> pool_; std::mutex m_; std::condition_variable cv_;
  std::deque<Job*> order_, todo_; bool done_ = false;
};

}  // namespace brbhost
