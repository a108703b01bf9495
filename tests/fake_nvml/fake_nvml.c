/*
 * fake_nvml.c -- a stand-in libnvidia-ml for tests and for BASELINE configs 1 and 5
 * (stubbed NVML: N fake GPUs x M MiB; scripted XID events). TEST INFRASTRUCTURE:
 * loaded only through gs_nvml_open(<path>) by tests/ and bench.py.
 *
 * Environment:
 *   FAKE_NVML_GPUS      number of GPUs (default 8)
 *   FAKE_NVML_MEM_MIB   total memory per GPU in MiB (default 16384)
 *   FAKE_NVML_MEM_FILE  path of a file holding that number instead; read at every memory query, so a
 *                       test can change what the GPUs report between two plugin restarts
 *   FAKE_NVML_NO_EVENTS comma list of GPU indices that answer NOT_SUPPORTED to
 *                       nvmlDeviceRegisterEvents
 *   FAKE_NVML_EVENT_FIFO path of a FIFO read for lines "<gpu_index_or_-1> <xid>": events for
 *                       a daemon running in another process
 * Events are injected with fake_nvml_push_event(gpu_index_or_-1, xid), an extra
 * export the tests call through ctypes, or through the FIFO above.
 * UUIDs follow SURVEY 8(d): GPU-%08x-0000-0000-0000-%012x from the index.
 */
#include <pthread.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

typedef int nvmlReturn_t;
typedef struct fake_dev { int index; } *nvmlDevice_t;
typedef struct fake_set { int dummy; } *nvmlEventSet_t;
#define NVML_SUCCESS 0
#define NVML_ERROR_UNINITIALIZED 1
#define NVML_ERROR_INVALID_ARGUMENT 2
#define NVML_ERROR_NOT_SUPPORTED 3
#define NVML_ERROR_INSUFFICIENT_SIZE 7
#define NVML_ERROR_TIMEOUT 10

typedef struct { unsigned int version; unsigned long long total, reserved, free, used; } nvmlMemory_v2_t;
typedef struct { unsigned long long total, free, used; } nvmlMemory_t;
typedef struct {
  char busIdLegacy[16]; unsigned int domain, bus, device, pciDeviceId, pciSubSystemId; char busId[32];
} nvmlPciInfo_t;
typedef struct {
  nvmlDevice_t device; unsigned long long eventType, eventData; unsigned int gpuInstanceId, computeInstanceId;
} nvmlEventData_t;

#define MAX_GPUS 64
static struct fake_dev g_devs[MAX_GPUS];
static struct fake_set g_set;
static int g_init = 0, g_n = 8;
static unsigned long long g_mem_mib = 16384;
static unsigned long long mem_mib_now(void) {
  const char *path = getenv("FAKE_NVML_MEM_FILE");
  if (path && *path) {
    FILE *f = fopen(path, "r");
    if (f) {
      unsigned long long v = 0;
      const int ok = fscanf(f, "%llu", &v) == 1;
      fclose(f);
      if (ok) return v;
    }
  }
  return g_mem_mib;
}
static int g_no_events[MAX_GPUS];

static pthread_mutex_t g_mu = PTHREAD_MUTEX_INITIALIZER;
static pthread_cond_t g_cv = PTHREAD_COND_INITIALIZER;
#define QCAP 256
static struct { int gpu; unsigned long long xid; unsigned long long type; } g_q[QCAP];
static int g_qh = 0, g_qt = 0;

#define EXPORT __attribute__((visibility("default")))

EXPORT void fake_nvml_push_event(int gpu, unsigned long long xid);

static char g_fifo[512];
static int g_fifo_started = 0;
static void *fifo_reader(void *arg) {
  (void)arg;
  for (;;) {
    FILE *f = fopen(g_fifo, "r"); /* blocks until a writer opens the FIFO */
    if (!f) return NULL;
    char line[128];
    while (fgets(line, sizeof line, f)) {
      int gpu;
      unsigned long long xid;
      if (sscanf(line, "%d %llu", &gpu, &xid) == 2) fake_nvml_push_event(gpu, xid);
    }
    fclose(f); /* writer went away: wait for the next one */
  }
}

EXPORT nvmlReturn_t nvmlInit_v2(void) {
  const char *s = getenv("FAKE_NVML_EVENT_FIFO");
  if (s && *s && !g_fifo_started) {
    pthread_t t;
    snprintf(g_fifo, sizeof g_fifo, "%s", s);
    g_fifo_started = 1;
    if (pthread_create(&t, NULL, fifo_reader, NULL) == 0) pthread_detach(t);
  }
  s = getenv("FAKE_NVML_GPUS");
  g_n = s ? atoi(s) : 8;
  if (g_n < 0) g_n = 0;
  if (g_n > MAX_GPUS) g_n = MAX_GPUS;
  s = getenv("FAKE_NVML_MEM_MIB");
  g_mem_mib = s ? strtoull(s, NULL, 10) : 16384ull;
  memset(g_no_events, 0, sizeof(g_no_events));
  s = getenv("FAKE_NVML_NO_EVENTS");
  if (s) {
    char buf[256];
    snprintf(buf, sizeof buf, "%s", s);
    for (char *tok = strtok(buf, ","); tok; tok = strtok(NULL, ",")) {
      int i = atoi(tok);
      if (i >= 0 && i < MAX_GPUS) g_no_events[i] = 1;
    }
  }
  for (int i = 0; i < MAX_GPUS; ++i) g_devs[i].index = i;
  g_qh = g_qt = 0;
  g_init = 1;
  return NVML_SUCCESS;
}
EXPORT nvmlReturn_t nvmlShutdown(void) { g_init = 0; return NVML_SUCCESS; }
EXPORT const char *nvmlErrorString(nvmlReturn_t r) {
  switch (r) {
    case NVML_SUCCESS: return "Success";
    case NVML_ERROR_UNINITIALIZED: return "Uninitialized";
    case NVML_ERROR_INVALID_ARGUMENT: return "Invalid Argument";
    case NVML_ERROR_NOT_SUPPORTED: return "Not Supported";
    case NVML_ERROR_TIMEOUT: return "Timeout";
    default: return "Unknown Error";
  }
}
EXPORT nvmlReturn_t nvmlDeviceGetCount_v2(unsigned int *n) {
  if (!g_init) return NVML_ERROR_UNINITIALIZED;
  *n = (unsigned)g_n;
  return NVML_SUCCESS;
}
EXPORT nvmlReturn_t nvmlDeviceGetHandleByIndex_v2(unsigned int i, nvmlDevice_t *d) {
  if (!g_init) return NVML_ERROR_UNINITIALIZED;
  if ((int)i >= g_n) return NVML_ERROR_INVALID_ARGUMENT;
  *d = &g_devs[i];
  return NVML_SUCCESS;
}
EXPORT nvmlReturn_t nvmlDeviceGetUUID(nvmlDevice_t d, char *buf, unsigned int len) {
  if (!g_init) return NVML_ERROR_UNINITIALIZED;
  if (len < 41) return NVML_ERROR_INSUFFICIENT_SIZE;
  snprintf(buf, len, "GPU-%08x-0000-0000-0000-%012x", d->index, d->index);
  return NVML_SUCCESS;
}
EXPORT nvmlReturn_t nvmlDeviceGetMemoryInfo_v2(nvmlDevice_t d, nvmlMemory_v2_t *m) {
  (void)d;
  if (!g_init) return NVML_ERROR_UNINITIALIZED;
  m->total = mem_mib_now() << 20;
  m->reserved = 0;
  m->free = m->total;
  m->used = 0;
  return NVML_SUCCESS;
}
EXPORT nvmlReturn_t nvmlDeviceGetMemoryInfo(nvmlDevice_t d, nvmlMemory_t *m) {
  (void)d;
  if (!g_init) return NVML_ERROR_UNINITIALIZED;
  m->total = mem_mib_now() << 20;
  m->free = m->total;
  m->used = 0;
  return NVML_SUCCESS;
}
EXPORT nvmlReturn_t nvmlDeviceGetPciInfo_v3(nvmlDevice_t d, nvmlPciInfo_t *p) {
  memset(p, 0, sizeof *p);
  p->bus = (unsigned)(0x10 + d->index);
  snprintf(p->busId, sizeof p->busId, "00000000:%02X:00.0", p->bus);
  return NVML_SUCCESS;
}
EXPORT nvmlReturn_t nvmlEventSetCreate(nvmlEventSet_t *s) {
  if (!g_init) return NVML_ERROR_UNINITIALIZED;
  *s = &g_set;
  return NVML_SUCCESS;
}
EXPORT nvmlReturn_t nvmlDeviceRegisterEvents(nvmlDevice_t d, unsigned long long types, nvmlEventSet_t s) {
  (void)types; (void)s;
  return g_no_events[d->index] ? NVML_ERROR_NOT_SUPPORTED : NVML_SUCCESS;
}
EXPORT nvmlReturn_t nvmlEventSetWait_v2(nvmlEventSet_t s, nvmlEventData_t *data, unsigned int timeoutms) {
  (void)s;
  struct timespec ts;
  clock_gettime(CLOCK_REALTIME, &ts);
  ts.tv_sec += timeoutms / 1000;
  ts.tv_nsec += (long)(timeoutms % 1000) * 1000000L;
  if (ts.tv_nsec >= 1000000000L) { ts.tv_sec++; ts.tv_nsec -= 1000000000L; }
  pthread_mutex_lock(&g_mu);
  while (g_qh == g_qt) {
    if (pthread_cond_timedwait(&g_cv, &g_mu, &ts) != 0) {
      pthread_mutex_unlock(&g_mu);
      return NVML_ERROR_TIMEOUT;
    }
  }
  int gpu = g_q[g_qh].gpu;
  data->device = (gpu >= 0 && gpu < g_n) ? &g_devs[gpu] : NULL;
  data->eventType = g_q[g_qh].type;
  data->eventData = g_q[g_qh].xid;
  data->gpuInstanceId = data->computeInstanceId = 0xFFFFFFFFu;
  g_qh = (g_qh + 1) % QCAP;
  pthread_mutex_unlock(&g_mu);
  return NVML_SUCCESS;
}
EXPORT nvmlReturn_t nvmlEventSetFree(nvmlEventSet_t s) { (void)s; return NVML_SUCCESS; }

/* test hook: gpu < 0 => event without a device identity ("empty UUID") */
EXPORT void fake_nvml_push_event(int gpu, unsigned long long xid) {
  pthread_mutex_lock(&g_mu);
  g_q[g_qt].gpu = gpu;
  g_q[g_qt].xid = xid;
  g_q[g_qt].type = 0x8ull; /* nvmlEventTypeXidCriticalError */
  g_qt = (g_qt + 1) % QCAP;
  pthread_cond_broadcast(&g_cv);
  pthread_mutex_unlock(&g_mu);
}
