// b200flow — shared-memory mailbox between a driver process and its worker group.
//
// Ray Train hands dataset shards to workers and receives `train.report(metrics,
// checkpoint)` messages back through the plasma object store / core worker (C++)
// — SURVEY.md §2E (containers/ray-pytorch/Dockerfile:1; the RayJob chart at
// charts/machine-learning/training/raytrain/templates/train.yaml:49-62).  On one box
// that hand-off is a POSIX shared-memory segment holding a single-producer /
// single-consumer byte ring per direction; this file is that ring.
//
// Layout of a segment:  Header | data[capacity].  Messages are length-prefixed
// (u32) and never wrap partially: if a message does not fit before the end of the
// ring a zero-length marker sends the reader back to offset 0.
#include <atomic>
#include <cerrno>
#include <cstdint>
#include <cstring>
#include <ctime>
#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

namespace {

struct Header {
  uint64_t magic;
  uint64_t capacity;
  std::atomic<uint64_t> head;   // total bytes written (monotonic)
  std::atomic<uint64_t> tail;   // total bytes consumed (monotonic)
  std::atomic<uint32_t> closed;
};

struct Channel {
  Header* hdr;
  uint8_t* data;
  size_t map_bytes;
  int owner;
  char name[128];
};

constexpr uint64_t MAGIC = 0xB200F10EC4A11E00ull;
constexpr uint32_t WRAP = 0xFFFFFFFFu;

void sleep_us(long us) {
  timespec ts{0, us * 1000};
  nanosleep(&ts, nullptr);
}

}  // namespace

extern "C" {

void* b2f_chan_create(const char* name, uint64_t capacity) {
  const size_t bytes = sizeof(Header) + capacity;
  shm_unlink(name);
  int fd = shm_open(name, O_CREAT | O_EXCL | O_RDWR, 0600);
  if (fd < 0) return nullptr;
  if (ftruncate(fd, static_cast<off_t>(bytes)) != 0) { close(fd); shm_unlink(name); return nullptr; }
  void* p = mmap(nullptr, bytes, PROT_READ | PROT_WRITE, MAP_SHARED, fd, 0);
  close(fd);
  if (p == MAP_FAILED) { shm_unlink(name); return nullptr; }
  Header* h = new (p) Header();
  h->capacity = capacity;
  h->head.store(0);
  h->tail.store(0);
  h->closed.store(0);
  h->magic = MAGIC;
  Channel* c = new Channel{h, reinterpret_cast<uint8_t*>(p) + sizeof(Header), bytes, 1, {0}};
  strncpy(c->name, name, sizeof(c->name) - 1);
  return c;
}

void* b2f_chan_open(const char* name) {
  int fd = shm_open(name, O_RDWR, 0600);
  if (fd < 0) return nullptr;
  struct stat st;
  if (fstat(fd, &st) != 0) { close(fd); return nullptr; }
  void* p = mmap(nullptr, static_cast<size_t>(st.st_size), PROT_READ | PROT_WRITE, MAP_SHARED, fd, 0);
  close(fd);
  if (p == MAP_FAILED) return nullptr;
  Header* h = reinterpret_cast<Header*>(p);
  if (h->magic != MAGIC) { munmap(p, static_cast<size_t>(st.st_size)); return nullptr; }
  Channel* c = new Channel{h, reinterpret_cast<uint8_t*>(p) + sizeof(Header),
                           static_cast<size_t>(st.st_size), 0, {0}};
  strncpy(c->name, name, sizeof(c->name) - 1);
  return c;
}

// Returns 0 on success, -1 if the message can never fit, -2 on timeout, -3 if closed.
int b2f_chan_send(void* ch, const void* buf, uint32_t len, long timeout_ms) {
  Channel* c = static_cast<Channel*>(ch);
  Header* h = c->hdr;
  const uint64_t cap = h->capacity;
  const uint64_t need = 4ull + len;
  if (need + 4 > cap) return -1;
  long waited_us = 0;
  for (;;) {
    if (h->closed.load()) return -3;
    const uint64_t head = h->head.load(std::memory_order_relaxed);
    const uint64_t tail = h->tail.load(std::memory_order_acquire);
    const uint64_t pos = head % cap;
    const uint64_t to_end = cap - pos;
    const uint64_t skip = to_end < need ? to_end : 0;  // wrap marker + padding
    if (cap - (head - tail) >= need + skip) {
      uint64_t w = head;
      if (skip) {
        if (to_end >= 4) memcpy(c->data + pos, &WRAP, 4);
        w += skip;
      }
      const uint64_t p2 = w % cap;
      memcpy(c->data + p2, &len, 4);
      memcpy(c->data + p2 + 4, buf, len);
      h->head.store(w + need, std::memory_order_release);
      return 0;
    }
    if (timeout_ms >= 0 && waited_us / 1000 >= timeout_ms) return -2;
    sleep_us(50);
    waited_us += 50;
  }
}

// Returns message length (copied into buf, up to buf_len), -2 on timeout, -3 if closed and
// drained, -4 if buf is too small (message left in place; required size in *need_out).
long long b2f_chan_recv(void* ch, void* buf, uint32_t buf_len, long timeout_ms, uint32_t* need_out) {
  Channel* c = static_cast<Channel*>(ch);
  Header* h = c->hdr;
  const uint64_t cap = h->capacity;
  long waited_us = 0;
  for (;;) {
    uint64_t tail = h->tail.load(std::memory_order_relaxed);
    const uint64_t head = h->head.load(std::memory_order_acquire);
    if (head != tail) {
      uint64_t pos = tail % cap;
      const uint64_t to_end = cap - pos;
      uint32_t len = WRAP;
      if (to_end >= 4) memcpy(&len, c->data + pos, 4);
      if (len == WRAP) {  // skip to the start of the ring
        tail += to_end;
        h->tail.store(tail, std::memory_order_release);
        continue;
      }
      if (len > buf_len) {
        if (need_out) *need_out = len;
        return -4;
      }
      memcpy(buf, c->data + pos + 4, len);
      h->tail.store(tail + 4ull + len, std::memory_order_release);
      return len;
    }
    if (h->closed.load()) return -3;
    if (timeout_ms >= 0 && waited_us / 1000 >= timeout_ms) return -2;
    sleep_us(50);
    waited_us += 50;
  }
}

void b2f_chan_close(void* ch) {
  Channel* c = static_cast<Channel*>(ch);
  c->hdr->closed.store(1);
}

void b2f_chan_free(void* ch) {
  Channel* c = static_cast<Channel*>(ch);
  const int owner = c->owner;
  char name[128];
  strncpy(name, c->name, sizeof(name));
  munmap(c->hdr, c->map_bytes);
  if (owner) shm_unlink(name);
  delete c;
}

}  // extern "C"
