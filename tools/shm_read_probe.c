// How should a restarted trainer pull a resident /dev/shm segment into its
// pinned bounce ring?  (1) memcpy out of a fresh mapping (page faults on first
// touch), (2) same after madvise(MADV_POPULATE_READ), (3) pread() from the fd
// (kernel copies from the page cache, no page tables).  CPU only.
// build: gcc -O2 -pthread tools/shm_read_probe.c -o /tmp/shm_read_probe
// usage: shm_read_probe GiB threads
#define _GNU_SOURCE
#include <fcntl.h>
#include <pthread.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/mman.h>
#include <time.h>
#include <unistd.h>

#define SLOT (8ull << 20)
static const char* NAME = "/dev/shm/fcp_shm_read_probe";
static uint64_t g_total;
static int g_fd, g_mode, g_threads;
static uint8_t* g_map;
static volatile uint64_t g_next;

static double now(void) {
  struct timespec t;
  clock_gettime(CLOCK_MONOTONIC, &t);
  return t.tv_sec + 1e-9 * t.tv_nsec;
}

static void* worker(void* arg) {
  (void)arg;
  uint8_t* buf = aligned_alloc(4096, SLOT);
  memset(buf, 1, SLOT);
  uint64_t sum = 0;
  for (;;) {
    uint64_t i = __sync_fetch_and_add(&g_next, 1);
    uint64_t off = i * SLOT;
    if (off >= g_total) break;
    uint64_t len = off + SLOT <= g_total ? SLOT : g_total - off;
    if (g_mode == 2) {
      uint64_t done = 0;
      while (done < len) {
        ssize_t r = pread(g_fd, buf + done, len - done, off + done);
        if (r <= 0) { perror("pread"); exit(1); }
        done += r;
      }
    } else {
      if (g_mode == 1) madvise(g_map + off, len, 22 /* MADV_POPULATE_READ */);
      memcpy(buf, g_map + off, len);
    }
    sum += buf[len - 1];
  }
  free(buf);
  return (void*)(uintptr_t)sum;
}

int main(int argc, char** argv) {
  g_total = (uint64_t)(atof(argc > 1 ? argv[1] : "4") * (1ull << 30));
  g_threads = argc > 2 ? atoi(argv[2]) : 8;
  g_fd = open(NAME, O_CREAT | O_RDWR | O_TRUNC, 0600);
  if (g_fd < 0 || ftruncate(g_fd, g_total)) { perror("create"); return 1; }
  uint8_t* w = mmap(NULL, g_total, PROT_READ | PROT_WRITE, MAP_SHARED | MAP_POPULATE, g_fd, 0);
  for (uint64_t o = 0; o < g_total; o += 4096) w[o] = (uint8_t)(o >> 12);
  munmap(w, g_total);
  const char* names[3] = {"memcpy from fresh mapping", "populate_read + memcpy", "pread"};
  for (int rep = 0; rep < 2; ++rep)
    for (g_mode = 0; g_mode < 3; ++g_mode) {
      g_map = mmap(NULL, g_total, PROT_READ, MAP_SHARED, g_fd, 0);  // fresh page tables
      g_next = 0;
      pthread_t th[64];
      double t0 = now();
      for (int i = 0; i < g_threads; ++i) pthread_create(&th[i], NULL, worker, NULL);
      for (int i = 0; i < g_threads; ++i) pthread_join(th[i], NULL);
      double dt = now() - t0;
      munmap(g_map, g_total);
      printf("%-28s threads=%d  %.3f s  %.1f GB/s\n", names[g_mode], g_threads, dt, g_total / 1e9 / dt);
    }
  close(g_fd);
  unlink(NAME);
  return 0;
}
