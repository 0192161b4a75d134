/*
 * CPU ORACLE (C part) — test infrastructure, NOT product code.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / `--impl reference` legs
 * may load liboracle.so.  Nothing under distributed-llm-dissemination_b200/ links or dlopens it.
 *
 *  1. oracle_fill / oracle_checksum: the same definitions as oracle/dissem_oracle.py
 *     (fill_bytes, checksum), in plain C so full-size layers are checked in seconds.
 *     The checksum is THIS REPO'S definition — the reference computes none (parity unpinned).
 *  2. oracle_tcp_run: a restatement of the reference's TCP byte-moving path, run over loopback
 *     on the host cores as the reported CPU baseline ("port": Go is not installed, so the
 *     reference itself cannot be built or run).  It follows, in /root/reference/distributor:
 *        transport.go:267-275   a FRESH connection is dialled for every layer message
 *        transport.go:308-331   JSON envelope {"type":2,"src":"<id>","payload":{SrcID,LayerID,
 *                               LayerSize,TotalSize,Offert}} written first
 *        transport.go:342-345   then raw bytes data[Offset:Offset+DataSize]
 *        transport.go:407-424   in writes of at most 256 KiB (the limiter's burst); oracle_tcp_run runs without
 *                               the limiter (rate = inf), oracle_tcp_run_ex restates the token bucket of
 *                               golang.org/x/time/rate v0.14.0 per transfer (module not under /root/reference;
 *                               published algorithm; PARITY UNPINNED: no reference test exercises it)
 *        transport.go:351-367   DiskLayer source (oracle_tcp_run_ex with a file): section of the file copied out
 *                               through a 32 KiB buffer (io.Copy's generic path), no limiter
 *        transport.go:97-103    receiver: streaming JSON decode of the envelope
 *        transport.go:201-207   fresh zeroed buffer of LayerSize, then read-full
 *        node.go:1354-1384      store, then ackMsg to the leader over the persistent control
 *                               connection (transport.go:228-256, :375-405)
 *        node.go:1520-1567      mode 3: count fragment bytes, ack when == TotalSize
 *        node.go:410-432        leader counts acks; completion when every assigned layer acked
 *        cmd/main.go:173-181    clock: all announces in -> last ack in
 *     One OS thread per in-flight message stands in for the goroutine the reference spawns
 *     (node.go:343, transport.go:90).
 */
#define _GNU_SOURCE
#include <arpa/inet.h>
#include <errno.h>
#include <fcntl.h>
#include <netinet/in.h>
#include <netinet/tcp.h>
#include <pthread.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/socket.h>
#include <time.h>
#include <unistd.h>

#define ORACLE_CHUNK (256 * 1024)
#define GOLDEN 0x9E3779B97F4A7C15ull

static inline uint64_t mix64(uint64_t z) {
    z ^= z >> 30; z *= 0xBF58476D1CE4E5B9ull;
    z ^= z >> 27; z *= 0x94D049BB133111EBull;
    z ^= z >> 31;
    return z;
}

/* kind: 0 zero (the reference's dummy layers, cmd/config.go:141,161), 1 pattern, 2 poison 0xA5 */
int oracle_fill(int kind, uint64_t layer_id, int64_t size, uint64_t seed, uint8_t *out) {
    if (size < 0) return -1;
    if (kind == 0) { memset(out, 0, (size_t)size); return 0; }
    if (kind == 2) { memset(out, 0xA5, (size_t)size); return 0; }
    if (kind != 1) return -1;
    uint64_t base = seed + (layer_id + 1) * GOLDEN;
    int64_t nfull = size / 8;
    for (int64_t i = 0; i < nfull; ++i) {
        uint64_t w = mix64(base + (uint64_t)i);
        memcpy(out + 8 * i, &w, 8);                 /* little-endian host assumed (x86-64) */
    }
    if (size % 8) {
        uint64_t w = mix64(base + (uint64_t)nfull);
        memcpy(out + 8 * nfull, &w, (size_t)(size % 8));
    }
    return 0;
}

/* out[2*c] = A_c, out[2*c+1] = B_c ; words little-endian u32, trailing partial word zero-extended;
 * i = word index within the layer.  See dissem_oracle.checksum. */
int oracle_checksum(const uint8_t *data, int64_t size, int64_t chunk_bytes, uint64_t *out) {
    if (size < 0 || chunk_bytes <= 0 || chunk_bytes % 4) return -1;
    int64_t n_chunks = (size + chunk_bytes - 1) / chunk_bytes;
    for (int64_t c = 0; c < n_chunks; ++c) {
        int64_t b0 = c * chunk_bytes, b1 = b0 + chunk_bytes;
        if (b1 > size) b1 = size;
        uint64_t A = 0, B = 0;
        int64_t p = b0;
        for (; p + 4 <= b1; p += 4) {
            uint32_t w;
            memcpy(&w, data + p, 4);
            A += w;
            B += ((uint64_t)(p >> 2) + 1) * (uint64_t)w;
        }
        if (p < b1) {
            uint32_t w = 0;
            memcpy(&w, data + p, (size_t)(b1 - p));
            A += w;
            B += ((uint64_t)(p >> 2) + 1) * (uint64_t)w;
        }
        out[2 * c] = A;
        out[2 * c + 1] = B;
    }
    return 0;
}

/* Bytes [off, off+len) of the layer oracle_fill would write (off a multiple of 8), for spot checks of full-size layers. */
int oracle_fill_range(int kind, uint64_t layer_id, int64_t off, int64_t len, uint64_t seed, uint8_t *out) {
    if (off < 0 || len < 0 || off % 8) return -1;
    if (kind == 0) { memset(out, 0, (size_t)len); return 0; }
    if (kind == 2) { memset(out, 0xA5, (size_t)len); return 0; }
    if (kind != 1) return -1;
    const uint64_t base = seed + (layer_id + 1) * GOLDEN;
    const int64_t w0 = off / 8, nfull = len / 8;
    for (int64_t i = 0; i < nfull; ++i) {
        uint64_t w = mix64(base + (uint64_t)(w0 + i));
        memcpy(out + 8 * i, &w, 8);
    }
    if (len % 8) {
        uint64_t w = mix64(base + (uint64_t)(w0 + nfull));
        memcpy(out + 8 * nfull, &w, (size_t)(len % 8));
    }
    return 0;
}

/* The table oracle_checksum would give for the bytes oracle_fill would write, without materialising the layer:
 * one chunk at a time through a chunk-sized buffer (full-size Llama-70B layers: 1.7 GB each, 80 of them).
 * Same two definitions, composed; tests/test_oracle_golden.py checks the composition against the two calls. */
int oracle_fill_checksum(int kind, uint64_t layer_id, int64_t size, uint64_t seed, int64_t chunk_bytes, uint64_t *out) {
    if (size < 0 || chunk_bytes <= 0 || chunk_bytes % 8 || kind < 0 || kind > 2) return -1;
    uint8_t *buf = (uint8_t *)malloc((size_t)chunk_bytes);
    if (!buf) return -1;
    const uint64_t base = seed + (layer_id + 1) * GOLDEN;
    int64_t n_chunks = (size + chunk_bytes - 1) / chunk_bytes;
    for (int64_t c = 0; c < n_chunks; ++c) {
        int64_t b0 = c * chunk_bytes, b1 = b0 + chunk_bytes;
        if (b1 > size) b1 = size;
        const int64_t len = b1 - b0;
        if (kind == 0) memset(buf, 0, (size_t)len);
        else if (kind == 2) memset(buf, 0xA5, (size_t)len);
        else {
            const int64_t w0 = b0 / 8, nfull = len / 8;       /* chunk_bytes % 8 == 0: chunks start on a u64 word */
            for (int64_t i = 0; i < nfull; ++i) {
                uint64_t w = mix64(base + (uint64_t)(w0 + i));
                memcpy(buf + 8 * i, &w, 8);
            }
            if (len % 8) {
                uint64_t w = mix64(base + (uint64_t)(w0 + nfull));
                memcpy(buf + 8 * nfull, &w, (size_t)(len % 8));
            }
        }
        uint64_t A = 0, B = 0;
        int64_t p = 0;
        for (; p + 4 <= len; p += 4) {
            uint32_t w;
            memcpy(&w, buf + p, 4);
            A += w;
            B += ((uint64_t)((b0 + p) >> 2) + 1) * (uint64_t)w;
        }
        if (p < len) {
            uint32_t w = 0;
            memcpy(&w, buf + p, (size_t)(len - p));
            A += w;
            B += ((uint64_t)((b0 + p) >> 2) + 1) * (uint64_t)w;
        }
        out[2 * c] = A;
        out[2 * c + 1] = B;
    }
    free(buf);
    return 0;
}

/* ------------------------------------------------------------------------------------------
 * loopback TCP restatement
 * ------------------------------------------------------------------------------------------ */
typedef struct {
    int32_t src, dst, layer, _pad;
    int64_t off, size, total;
} oracle_xfer;

typedef struct run_s run_t;

typedef struct {
    run_t *run;
    int node;
    int fd;            /* listener */
    pthread_t thr;
} listener_t;

typedef struct {
    int layer;
    int64_t total, got;
    uint8_t *buf;      /* assembled layer at the destination (kept for verification) */
} recv_layer_t;

struct run_s {
    int n_nodes, leader, base_port;
    const oracle_xfer *xf;
    int n_xf;
    const uint8_t *const *src_base;      /* per transfer: base pointer of the SOURCE layer (NULL: read it from src_path) */
    const char *const *src_path;         /* per transfer: file holding the source layer (DiskLayer), or NULL (may be NULL itself) */
    const int64_t *rate;                 /* per transfer: LimitRate in bytes/s, <= 0 = no limiter (may be NULL) */
    listener_t *lis;
    pthread_mutex_t mu;
    pthread_cond_t cv;
    int acks, acks_needed;
    int failed;
    /* per (dst, layer) assembly state */
    recv_layer_t *rl;
    int *rl_node;
    int n_rl;
    int ctrl_fd[64];                     /* node -> persistent control connection to the leader */
    pthread_mutex_t ctrl_mu[64];
    volatile int stop;
    int live_conns;                      /* connection threads still running */
};

static double now_s(void) {
    struct timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return ts.tv_sec + 1e-9 * ts.tv_nsec;
}

static int write_full(int fd, const void *p, size_t n) {
    const uint8_t *b = (const uint8_t *)p;
    while (n) {
        ssize_t w = send(fd, b, n, MSG_NOSIGNAL);
        if (w < 0) { if (errno == EINTR) continue; return -1; }
        b += w; n -= (size_t)w;
    }
    return 0;
}

static int read_full(int fd, void *p, size_t n) {
    uint8_t *b = (uint8_t *)p;
    while (n) {
        ssize_t r = recv(fd, b, n, 0);
        if (r == 0) return -1;
        if (r < 0) { if (errno == EINTR) continue; return -1; }
        b += r; n -= (size_t)r;
    }
    return 0;
}

static int dial(int port) {
    int fd = socket(AF_INET, SOCK_STREAM, 0);
    if (fd < 0) return -1;
    struct sockaddr_in a;
    memset(&a, 0, sizeof a);
    a.sin_family = AF_INET;
    a.sin_port = htons((uint16_t)port);
    a.sin_addr.s_addr = htonl(INADDR_LOOPBACK);
    if (connect(fd, (struct sockaddr *)&a, sizeof a) < 0) { close(fd); return -1; }
    return fd;
}

static void fail(run_t *r) {
    pthread_mutex_lock(&r->mu);
    r->failed = 1;
    pthread_cond_broadcast(&r->cv);
    pthread_mutex_unlock(&r->mu);
}

/* ack: {"type":1,"src":"<id>","payload":{"SrcID":<id>,"LayerID":<layer>}}  (message.go:62-74; `location`
 * is unexported and therefore absent on the wire) */
static int send_ack(run_t *r, int node, int layer) {
    char msg[160];
    int n = snprintf(msg, sizeof msg, "{\"type\":1,\"src\":\"%d\",\"payload\":{\"SrcID\":%d,\"LayerID\":%d}}",
                     node, node, layer);
    if (node == r->leader) {            /* a send to one's own address short-circuits, transport.go:282-286 */
        pthread_mutex_lock(&r->mu);
        r->acks++;
        pthread_cond_broadcast(&r->cv);
        pthread_mutex_unlock(&r->mu);
        return 0;
    }
    pthread_mutex_lock(&r->ctrl_mu[node]);          /* per-connection mutex, transport.go:376-377 */
    if (r->ctrl_fd[node] < 0) r->ctrl_fd[node] = dial(r->base_port + r->leader);   /* lazy, :228-256 */
    int rc = r->ctrl_fd[node] < 0 ? -1 : write_full(r->ctrl_fd[node], msg, (size_t)n);
    pthread_mutex_unlock(&r->ctrl_mu[node]);
    return rc;
}

static long json_int(const char *s, const char *key) {
    const char *p = strstr(s, key);
    if (!p) return -1;
    p += strlen(key);
    while (*p == '"' || *p == ':' || *p == ' ') ++p;
    return strtol(p, NULL, 10);
}

typedef struct { run_t *run; int node; int fd; } conn_t;

/* handleIncomingMsg, transport.go:97-225: decode concatenated JSON envelopes; a type-2 envelope is
 * followed by exactly LayerSize raw bytes. */
static void *conn_main(void *arg) {
    conn_t *c = (conn_t *)arg;
    run_t *r = c->run;
    char buf[8192];
    size_t have = 0;
    for (;;) {
        /* find one complete JSON object in buf[0..have) by brace depth (no braces occur in strings here) */
        size_t end = 0;
        int depth = 0, started = 0;
        for (size_t i = 0; i < have; ++i) {
            if (buf[i] == '{') { depth++; started = 1; }
            else if (buf[i] == '}') { depth--; if (started && depth == 0) { end = i + 1; break; } }
        }
        if (!end) {
            if (have == sizeof buf - 1) { fail(r); break; }
            ssize_t n = recv(c->fd, buf + have, sizeof buf - 1 - have, 0);
            if (n <= 0) break;                       /* peer closed: io.EOF, transport.go:106-111 */
            have += (size_t)n;
            continue;
        }
        char hdr[1024];
        size_t hl = end < sizeof hdr - 1 ? end : sizeof hdr - 1;
        memcpy(hdr, buf, hl);
        hdr[hl] = 0;
        long type = json_int(hdr, "\"type\"");
        if (type == 2) {
            long layer = json_int(hdr, "\"LayerID\"");
            long long lsize = strtoll(strstr(hdr, "\"LayerSize\":") + 12, NULL, 10);
            long long total = strtoll(strstr(hdr, "\"TotalSize\":") + 12, NULL, 10);
            long long off = strtoll(strstr(hdr, "\"Offert\":") + 9, NULL, 10);
            /* buf = make(LayerData, temp.LayerSize): fresh zeroed memory, transport.go:201 */
            uint8_t *body = (uint8_t *)calloc((size_t)(lsize > 0 ? lsize : 1), 1);
            if (!body) { fail(r); break; }
            size_t pre = have - end;                 /* bytes the decoder had buffered, :202 */
            if ((long long)pre > lsize) pre = (size_t)lsize;
            memcpy(body, buf + end, pre);
            if (read_full(c->fd, body + pre, (size_t)lsize - pre) < 0) { free(body); fail(r); break; }
            have = 0;
            /* handleLayerMsg (node.go:1354-1384 / :1520-1567) */
            int do_ack = 0;
            pthread_mutex_lock(&r->mu);
            recv_layer_t *rl = NULL;
            for (int i = 0; i < r->n_rl; ++i)
                if (r->rl_node[i] == c->node && r->rl[i].layer == (int)layer) { rl = &r->rl[i]; break; }
            pthread_mutex_unlock(&r->mu);
            if (!rl) { free(body); fail(r); break; }
            if (lsize == total && off == 0) {        /* whole layer: the received buffer IS the layer */
                pthread_mutex_lock(&r->mu);
                free(rl->buf);
                rl->buf = body;
                rl->got = total;
                do_ack = 1;
                pthread_mutex_unlock(&r->mu);
            } else {                                  /* fragment: place at its offset (node.go:1545-1547 intent) */
                pthread_mutex_lock(&r->mu);
                if (!rl->buf) rl->buf = (uint8_t *)calloc((size_t)(total > 0 ? total : 1), 1);
                uint8_t *dstp = rl->buf;
                pthread_mutex_unlock(&r->mu);
                if (!dstp) { free(body); fail(r); break; }
                memcpy(dstp + off, body, (size_t)lsize);
                free(body);
                pthread_mutex_lock(&r->mu);
                rl->got += lsize;
                do_ack = (rl->got == total);
                pthread_mutex_unlock(&r->mu);
            }
            if (do_ack && send_ack(r, c->node, (int)layer) < 0) { fail(r); break; }
        } else {
            if (type == 1) {                          /* handleAckMsg, node.go:410-432 */
                pthread_mutex_lock(&r->mu);
                r->acks++;
                pthread_cond_broadcast(&r->cv);
                pthread_mutex_unlock(&r->mu);
            }
            memmove(buf, buf + end, have - end);
            have -= end;
        }
    }
    close(c->fd);
    free(c);
    __atomic_fetch_sub(&r->live_conns, 1, __ATOMIC_SEQ_CST);   /* last touch of *r */
    return NULL;
}

static void *listener_main(void *arg) {
    listener_t *l = (listener_t *)arg;
    for (;;) {
        int fd = accept(l->fd, NULL, NULL);
        if (fd < 0) {
            if (errno == EINTR) continue;
            return NULL;                              /* listener closed */
        }
        conn_t *c = (conn_t *)malloc(sizeof *c);
        c->run = l->run; c->node = l->node; c->fd = fd;
        pthread_t t;                                  /* go t.handleIncomingMsg(conn), transport.go:90 */
        __atomic_fetch_add(&l->run->live_conns, 1, __ATOMIC_SEQ_CST);
        if (pthread_create(&t, NULL, conn_main, c) != 0) {
            __atomic_fetch_sub(&l->run->live_conns, 1, __ATOMIC_SEQ_CST);
            close(fd); free(c); fail(l->run); return NULL;
        }
        pthread_detach(t);
    }
}

typedef struct { run_t *run; int i; } send_arg_t;

/* TcpTransport.Send layer branch + sendLayerMsg + writeWithLimit, transport.go:267-275, :308-373, :407-424 */
static void *send_main(void *arg) {
    send_arg_t *sa = (send_arg_t *)arg;
    run_t *r = sa->run;
    const oracle_xfer *x = &r->xf[sa->i];
    int fd = dial(r->base_port + x->dst);
    if (fd < 0) { fail(r); free(sa); return NULL; }
    char hdr[256];
    int n = snprintf(hdr, sizeof hdr,
                     "{\"type\":2,\"src\":\"%d\",\"payload\":{\"SrcID\":%d,\"LayerID\":%d,\"LayerSize\":%lld,"
                     "\"TotalSize\":%lld,\"Offert\":%lld}}",
                     x->src, x->src, x->layer, (long long)x->size, (long long)x->total, (long long)x->off);
    int rc = write_full(fd, hdr, (size_t)n);
    const char *path = r->src_path ? r->src_path[sa->i] : NULL;
    if (path) {
        /* DiskLayer, transport.go:351-367: io.Copy(conn, io.NewSectionReader(f, Offset, DataSize)).  A SectionReader is
         * not an *os.File, so TCPConn.ReadFrom falls back to the generic copy loop: read into a 32 KiB buffer, write it
         * out (the comment at :362 hopes for sendfile; the code path is this one).  No limiter on this branch. */
        int ffd = open(path, O_RDONLY);
        if (ffd < 0) rc = -1;
        static __thread uint8_t cp[32 * 1024];
        int64_t pos = x->off, left = x->size;
        while (rc == 0 && left > 0) {
            ssize_t k = pread(ffd, cp, (size_t)(left < (int64_t)sizeof cp ? left : (int64_t)sizeof cp), (off_t)pos);
            if (k < 0 && errno == EINTR) continue;
            if (k <= 0) { rc = -1; break; }
            rc = write_full(fd, cp, (size_t)k);
            pos += k; left -= k;
        }
        if (ffd >= 0) close(ffd);
    } else {
        /* writeWithLimit, transport.go:407-424: limiter := rate.NewLimiter(rate.Limit(limitRate), 256 KiB); per write
         * n = min(len(data), burst); limiter.WaitN(n).  golang.org/x/time/rate v0.14.0 (go.mod:8), restated: the bucket
         * starts full; reserveN advances it by elapsed*limit (capped at burst), takes n, and a negative balance waits
         * -tokens/limit seconds.  rate <= 0 here means "no limiter" (with limit 0 the reference never finishes). */
        const double limit = (r->rate && r->rate[sa->i] > 0) ? (double)r->rate[sa->i] : 0.0;
        double tokens = (double)ORACLE_CHUNK, last = now_s();
        const uint8_t *p = r->src_base[sa->i] + x->off;
        int64_t left = x->size;
        while (rc == 0 && left > 0) {
            int64_t k = left < ORACLE_CHUNK ? left : ORACLE_CHUNK;   /* min(len(data), limiter.Burst()) */
            if (limit > 0.0) {
                double now = now_s();
                tokens += (now - last) * limit;
                if (tokens > (double)ORACLE_CHUNK) tokens = (double)ORACLE_CHUNK;
                tokens -= (double)k;
                last = now;
                if (tokens < 0.0) {
                    double wait = -tokens / limit;
                    struct timespec ts;
                    ts.tv_sec = (time_t)wait;
                    ts.tv_nsec = (long)((wait - (double)ts.tv_sec) * 1e9);
                    while (nanosleep(&ts, &ts) != 0 && errno == EINTR) {}
                }
            }
            rc = write_full(fd, p, (size_t)k);
            p += k; left -= k;
        }
    }
    close(fd);                                                   /* defer conn.Close(), :273 */
    if (rc < 0) fail(r);
    free(sa);
    return NULL;
}

/* Runs the transfer list once.  Returns 0 on success; *seconds = completion time (clock of
 * cmd/main.go:173-181); if verify != 0 every destination layer is compared with its source bytes and
 * csum_out (may be NULL) receives, per transfer, A and B of chunk 0 of the destination layer as a spot
 * check.  max_inflight > 0 bounds concurrently running sender threads (0 = one per transfer, as the
 * reference's goroutine-per-send). */
int oracle_tcp_run_ex(int n_nodes, int leader, int base_port, const oracle_xfer *xf, int n_xf,
                      const uint8_t *const *src_base, const char *const *src_path, const int64_t *rate, int verify,
                      double *seconds, int64_t *bytes_moved);

int oracle_tcp_run(int n_nodes, int leader, int base_port, const oracle_xfer *xf, int n_xf,
                   const uint8_t *const *src_base, int verify, double *seconds, int64_t *bytes_moved) {
    return oracle_tcp_run_ex(n_nodes, leader, base_port, xf, n_xf, src_base, NULL, NULL, verify, seconds, bytes_moved);
}

/* The same with, per transfer, an optional source FILE (the DiskLayer branch of sendLayerMsg) and an optional LimitRate
 * (writeWithLimit's token bucket).  With a file source, src_base[i] may still be given: it is then what verification
 * compares the destination with (the expected contents of the file). */
int oracle_tcp_run_ex(int n_nodes, int leader, int base_port, const oracle_xfer *xf, int n_xf,
                      const uint8_t *const *src_base, const char *const *src_path, const int64_t *rate, int verify,
                      double *seconds, int64_t *bytes_moved) {
    if (n_nodes <= 0 || n_nodes > 64 || n_xf < 0) return -1;
    run_t r;
    memset(&r, 0, sizeof r);
    r.n_nodes = n_nodes; r.leader = leader; r.base_port = base_port;
    r.xf = xf; r.n_xf = n_xf; r.src_base = src_base; r.src_path = src_path; r.rate = rate;
    pthread_mutex_init(&r.mu, NULL);
    pthread_cond_init(&r.cv, NULL);
    for (int i = 0; i < 64; ++i) { r.ctrl_fd[i] = -1; pthread_mutex_init(&r.ctrl_mu[i], NULL); }
    /* distinct (dst, layer) pairs = acks the leader waits for */
    r.rl = (recv_layer_t *)calloc((size_t)(n_xf ? n_xf : 1), sizeof *r.rl);
    r.rl_node = (int *)calloc((size_t)(n_xf ? n_xf : 1), sizeof(int));
    for (int i = 0; i < n_xf; ++i) {
        int found = 0;
        for (int j = 0; j < r.n_rl; ++j)
            if (r.rl_node[j] == xf[i].dst && r.rl[j].layer == xf[i].layer) { found = 1; break; }
        if (!found) {
            r.rl_node[r.n_rl] = xf[i].dst;
            r.rl[r.n_rl].layer = xf[i].layer;
            r.rl[r.n_rl].total = xf[i].total;
            r.n_rl++;
        }
    }
    r.acks_needed = r.n_rl;
    r.lis = (listener_t *)calloc((size_t)n_nodes, sizeof *r.lis);
    int ok = 1;
    for (int i = 0; i < n_nodes && ok; ++i) {
        int fd = socket(AF_INET, SOCK_STREAM, 0);
        int one = 1;
        setsockopt(fd, SOL_SOCKET, SO_REUSEADDR, &one, sizeof one);
        struct sockaddr_in a;
        memset(&a, 0, sizeof a);
        a.sin_family = AF_INET;
        a.sin_port = htons((uint16_t)(base_port + i));
        a.sin_addr.s_addr = htonl(INADDR_LOOPBACK);
        if (fd < 0 || bind(fd, (struct sockaddr *)&a, sizeof a) < 0 || listen(fd, 256) < 0) { ok = 0; r.lis[i].fd = fd; break; }
        r.lis[i].run = &r; r.lis[i].node = i; r.lis[i].fd = fd;
        pthread_create(&r.lis[i].thr, NULL, listener_main, &r.lis[i]);
    }
    int rc = -1;
    if (ok) {
        double t0 = now_s();                          /* <-leader.StartDistribution() */
        pthread_t *st = (pthread_t *)calloc((size_t)(n_xf ? n_xf : 1), sizeof *st);
        for (int i = 0; i < n_xf; ++i) {              /* go sendLayer(...), node.go:343-349 */
            send_arg_t *sa = (send_arg_t *)malloc(sizeof *sa);
            sa->run = &r; sa->i = i;
            if (pthread_create(&st[i], NULL, send_main, sa) != 0) { free(sa); fail(&r); st[i] = 0; }
        }
        pthread_mutex_lock(&r.mu);
        while (r.acks < r.acks_needed && !r.failed) pthread_cond_wait(&r.cv, &r.mu);
        pthread_mutex_unlock(&r.mu);
        double t1 = now_s();                          /* <-leader.Ready() */
        for (int i = 0; i < n_xf; ++i) if (st[i]) pthread_join(st[i], NULL);
        free(st);
        if (!r.failed) {
            rc = 0;
            if (seconds) *seconds = t1 - t0;
            int64_t b = 0;
            for (int i = 0; i < n_xf; ++i) b += xf[i].size;
            if (bytes_moved) *bytes_moved = b;
            if (verify) {
                for (int i = 0; i < n_xf && rc == 0; ++i) {
                    recv_layer_t *rl = NULL;
                    for (int j = 0; j < r.n_rl; ++j)
                        if (r.rl_node[j] == xf[i].dst && r.rl[j].layer == xf[i].layer) rl = &r.rl[j];
                    if (!rl || !rl->buf || rl->got != rl->total) { rc = -2; break; }
                    if (src_base[i] && memcmp(rl->buf + xf[i].off, src_base[i] + xf[i].off, (size_t)xf[i].size) != 0) rc = -2;
                }
            }
        }
    }
    for (int i = 0; i < n_nodes; ++i) {
        if (r.lis[i].fd > 0) { shutdown(r.lis[i].fd, SHUT_RDWR); close(r.lis[i].fd); }
        if (r.lis[i].thr) pthread_join(r.lis[i].thr, NULL);
    }
    for (int i = 0; i < 64; ++i) if (r.ctrl_fd[i] >= 0) close(r.ctrl_fd[i]);
    for (int spin = 0; spin < 5000 && __atomic_load_n(&r.live_conns, __ATOMIC_SEQ_CST) > 0; ++spin)
        usleep(1000);                                 /* detached connection threads see EOF and exit */
    for (int i = 0; i < r.n_rl; ++i) free(r.rl[i].buf);
    free(r.rl); free(r.rl_node); free(r.lis);
    return rc;
}
