// output.cu -- the reference's field writer (src/serial/io.c:61-120: `output()`, called every n_print-th step with
// prefix out/<n>, main.c:138-143) at GPU speed.
//
// Format, byte for byte (io.c:89-112): three header lines t, a, b as "%.5f\n"; then one text row per grid row
// j = 0 .. j_max+1 with one "%.5f " (trailing blank) per column i = 0 .. i_max+1; `_u.txt` omits column i_max+1,
// `_v.txt` omits the VALUES of row j_max+1 but keeps its newline.
//
// The reference formats ~200 M values through fprintf on the thread that also runs the solver.  Here:
//   1. the fields are snapshotted on the device (three device-to-device copies on the solver's stream, ~1 ms at
//      8192^2) so the time loop can go on at once;
//   2. the snapshot travels to pinned host memory on a SIDE stream (x-fastest rows = the file's row order, no transpose);
//   3. a background host thread waits for the copy and formats row blocks on all host cores -- "%.5f" is done in
//      exact integer arithmetic (a binary64 times 10^5 is at most 70 bits; round half to even on the exact value, which
//      is what glibc's printf does), so the bytes equal fprintf's -- and writes each file with one stream of large
//      write()s.
// nsb_output() is the synchronous form (returns when the files are closed), nsb_output_async() / nsb_output_wait()
// the overlapped one; nsb_group_output*() write the same three files for a y-slab decomposition (every slab
// contributes its rows, in rank order).
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <atomic>
#include <string>
#include <thread>
#include <vector>

#include "nsb_internal.h"

// ------------------------------------------------------------------ "%.5f" in integer arithmetic
// Writes printf("%.5f", x) at `o`, returns the number of bytes.  Exact for every finite |x| < 2^46 (7e13);
// beyond that, and for Inf / NaN, it defers to snprintf (never taken for field data).
static inline int fmt5(double x, char* o) {
    unsigned long long b;
    memcpy(&b, &x, 8);
    const int neg = (int)(b >> 63);
    const int ex = (int)((b >> 52) & 0x7ff);
    unsigned long long m = b & 0xfffffffffffffull;
    if (ex == 0x7ff || ex >= 1023 + 46) return snprintf(o, 400, "%.5f", x);
    char* p = o;
    if (neg) *p++ = '-';
    int e;                                                     // x = m * 2^e
    if (ex == 0) e = -1074; else { m |= 1ull << 52; e = ex - 1075; }
    unsigned long long q;                                      // round_half_even(|x| * 10^5), < 2^63
    {
        const unsigned __int128 N = (unsigned __int128)m * 100000u;   // < 2^70
        const int s = -e;                                      // > 0 here (ex < 1069)
        if (s >= 100) q = 0;                                   // |x| * 10^5 < 2^-30: rounds to 0, no tie possible
        else {
            q = (unsigned long long)(N >> s);
            const unsigned __int128 rem = N & ((((unsigned __int128)1) << s) - 1);
            const unsigned __int128 half = ((unsigned __int128)1) << (s - 1);
            if (rem > half || (rem == half && (q & 1))) q++;
        }
    }
    unsigned long long ip = q / 100000u;
    unsigned int fr = (unsigned int)(q % 100000u);
    char tmp[24];
    int n = 0;
    do { tmp[n++] = (char)('0' + ip % 10); ip /= 10; } while (ip);
    while (n) *p++ = tmp[--n];
    *p++ = '.';
    p[4] = (char)('0' + fr % 10); fr /= 10;
    p[3] = (char)('0' + fr % 10); fr /= 10;
    p[2] = (char)('0' + fr % 10); fr /= 10;
    p[1] = (char)('0' + fr % 10); fr /= 10;
    p[0] = (char)('0' + fr);
    return (int)(p + 5 - o);
}

extern "C" int nsb_debug_fmt5(double x, char* out) { int n = fmt5(x, out); out[n] = 0; return n; }

// ------------------------------------------------------------------ the job
struct OutSlab {
    nsb_ctx* c;
    int l0, l1;                 // local rows this slab contributes (inclusive)
};

struct OutState {               // lives in the first context of the job (nsb_ctx::out_state)
    std::thread worker;
    bool running = false;
    int status = NSB_OK;
    char err[256] = "";
};

struct OutBuffers {             // per context (nsb_ctx::out_buffers): snapshot grids, pinned host copies, side stream
    double* d_snap[3] = {NULL, NULL, NULL};
    double* h[3] = {NULL, NULL, NULL};
    cudaStream_t side = NULL;
    cudaEvent_t snapped = NULL, landed = NULL;
};

static OutBuffers* out_buffers(nsb_ctx* c, bool with_snapshot) {
    OutBuffers* ob = (OutBuffers*)c->out_buffers;
    if (!ob) { ob = new OutBuffers(); c->out_buffers = ob; }
    const size_t dense = ((size_t)c->g.ni + 2) * ((size_t)c->g.nj + 2) * sizeof(double);
    for (int k = 0; k < 3; k++) {
        if (!ob->h[k] && cudaHostAlloc((void**)&ob->h[k], dense, cudaHostAllocDefault) != cudaSuccess) return NULL;
        if (with_snapshot && !ob->d_snap[k] && cudaMalloc((void**)&ob->d_snap[k], c->g.elems * sizeof(double)) != cudaSuccess) return NULL;
    }
    if (!ob->side && cudaStreamCreateWithFlags(&ob->side, cudaStreamNonBlocking) != cudaSuccess) return NULL;
    if (!ob->snapped && cudaEventCreateWithFlags(&ob->snapped, cudaEventDisableTiming) != cudaSuccess) return NULL;
    if (!ob->landed && cudaEventCreateWithFlags(&ob->landed, cudaEventDisableTiming) != cudaSuccess) return NULL;
    return ob;
}

void nsb_out_release(nsb_ctx* c) {                             // nsb_destroy
    OutState* st = (OutState*)c->out_state;
    if (st) {
        if (st->worker.joinable()) st->worker.join();
        delete st;
        c->out_state = NULL;
    }
    OutBuffers* ob = (OutBuffers*)c->out_buffers;
    if (ob) {
        for (int k = 0; k < 3; k++) { if (ob->d_snap[k]) cudaFree(ob->d_snap[k]); if (ob->h[k]) cudaFreeHost(ob->h[k]); }
        if (ob->side) cudaStreamDestroy(ob->side);
        if (ob->snapped) cudaEventDestroy(ob->snapped);
        if (ob->landed) cudaEventDestroy(ob->landed);
        delete ob;
        c->out_buffers = NULL;
    }
}

static int n_format_threads() {
    int n = (int)std::thread::hardware_concurrency();
    const char* env = getenv("NSB_OUTPUT_THREADS");
    if (env && atoi(env) > 0) n = atoi(env);
    if (n < 1) n = 1;
    if (n > 64) n = 64;
    return n;
}

// Host part: every slab's fields are in its pinned buffers.  Formats row blocks on all cores, writes in order.
static int out_format_and_write(const std::vector<OutSlab>& slabs, double t, double a, double b, const std::string& prefix,
                                char* err, size_t errlen) {
    const char* suffix[3] = {"_u.txt", "_v.txt", "_p.txt"};
    FILE* fp[3];
    for (int k = 0; k < 3; k++) fp[k] = fopen((prefix + suffix[k]).c_str(), "w");
    if (!fp[0] || !fp[1] || !fp[2]) {
        for (int k = 0; k < 3; k++) if (fp[k]) fclose(fp[k]);
        snprintf(err, errlen, "Error opening one ore more output files. Make sure the directory 'out' exists.");   // io.c:85
        return NSB_ERR_IO;
    }
    const nsb_ctx* c0 = slabs[0].c;
    const size_t nx = (size_t)c0->g.ni + 2;
    const int J = c0->J;
    // row blocks over all slabs
    struct Block { int slab, l0, l1; };
    std::vector<Block> blocks;
    const int rows_per_block = 32;
    for (size_t s = 0; s < slabs.size(); s++)
        for (int l = slabs[s].l0; l <= slabs[s].l1; l += rows_per_block)
            blocks.push_back({(int)s, l, l + rows_per_block - 1 < slabs[s].l1 ? l + rows_per_block - 1 : slabs[s].l1});
    int rc = NSB_OK;
    for (int k = 0; k < 3 && rc == NSB_OK; k++) {
        {
            char hdr[1300];
            int n = 0;
            n += fmt5(t, hdr + n); hdr[n++] = '\n';                 // io.c:89-99
            n += fmt5(a, hdr + n); hdr[n++] = '\n';
            n += fmt5(b, hdr + n); hdr[n++] = '\n';
            fwrite(hdr, 1, (size_t)n, fp[k]);
        }
        const size_t ncol = (k == 0) ? nx - 1 : nx;             // u omits i = i_max+1
        std::vector<std::string> text(blocks.size());
        std::atomic<size_t> next(0);
        auto work = [&]() {
            // a value takes at most 22 bytes on the integer path; the snprintf path (|x| >= 2^46, Inf, NaN) up to ~320:
            // the buffer is flushed whenever fewer than 512 bytes are left
            std::vector<char> buf((size_t)rows_per_block * (ncol * 24 + 2) + 1024);
            char* const end = buf.data() + buf.size();
            for (;;) {
                const size_t bi = next.fetch_add(1);
                if (bi >= blocks.size()) break;
                const Block& B = blocks[bi];
                const nsb_ctx* c = slabs[B.slab].c;
                const double* h = ((const OutBuffers*)c->out_buffers)->h[k];
                char* p = buf.data();
                text[bi].clear();
                for (int l = B.l0; l <= B.l1; l++) {
                    const int jg = c->j_lo + l;                  // global row
                    if (!(k == 1 && jg == J + 1)) {              // v omits the values of row j_max+1 ...
                        const double* row = h + (size_t)l * nx;
                        for (size_t i = 0; i < ncol; i++) {
                            if (end - p < 512) { text[bi].append(buf.data(), (size_t)(p - buf.data())); p = buf.data(); }
                            p += fmt5(row[i], p); *p++ = ' ';
                        }
                    }
                    *p++ = '\n';                                 // ... but keeps its newline
                }
                text[bi].append(buf.data(), (size_t)(p - buf.data()));
            }
        };
        const int nt = n_format_threads();
        std::vector<std::thread> pool;
        for (int q = 1; q < nt; q++) pool.emplace_back(work);
        work();
        for (auto& th : pool) th.join();
        for (size_t bi = 0; bi < blocks.size(); bi++)
            if (fwrite(text[bi].data(), 1, text[bi].size(), fp[k]) != text[bi].size()) { rc = NSB_ERR_IO; break; }
    }
    for (int k = 0; k < 3; k++) if (fclose(fp[k]) != 0) rc = NSB_ERR_IO;
    if (rc != NSB_OK) snprintf(err, errlen, "writing %s_*.txt failed", prefix.c_str());
    return rc;
}

#define OCU(c, call)                                                                                          \
    do {                                                                                                      \
        cudaError_t e__ = (call);                                                                             \
        if (e__ != cudaSuccess) return nsb_fail((c), NSB_ERR_CUDA, "%s failed: %s", #call, cudaGetErrorString(e__)); \
    } while (0)

// rows a slab contributes to the global files
static OutSlab out_rows(nsb_ctx* c) {
    OutSlab s;
    s.c = c;
    s.l0 = c->wall_bottom ? 0 : 1;
    s.l1 = c->wall_top ? c->g.nj + 1 : c->g.nj;
    return s;
}

// Enqueue the device side of a dump for one context: [snapshot on the solver's stream,] copies to pinned memory.
static int out_enqueue(nsb_ctx* c, bool async) {
    OCU(c, cudaSetDevice(c->device));
    OutBuffers* ob = out_buffers(c, async);
    if (!ob) return nsb_fail(c, NSB_ERR_CUDA, "output: cannot allocate the staging buffers");
    const double* src[3] = {c->u, c->v, c->pbuf[c->pcur]};
    const size_t nx = (size_t)c->g.ni + 2, ny = (size_t)c->g.nj + 2;
    cudaStream_t copy_stream = c->stream;
    if (async) {
        for (int k = 0; k < 3; k++)
            OCU(c, cudaMemcpyAsync(ob->d_snap[k], src[k], c->g.elems * sizeof(double), cudaMemcpyDeviceToDevice, c->stream));
        OCU(c, cudaEventRecord(ob->snapped, c->stream));
        OCU(c, cudaStreamWaitEvent(ob->side, ob->snapped, 0));
        for (int k = 0; k < 3; k++) src[k] = ob->d_snap[k];
        copy_stream = ob->side;
    }
    for (int k = 0; k < 3; k++)
        OCU(c, cudaMemcpy2DAsync(ob->h[k], nx * sizeof(double), src[k] + (size_t)NSB_ROW0 * c->g.pitch + NSB_COL0,
                                 (size_t)c->g.pitch * sizeof(double), nx * sizeof(double), ny, cudaMemcpyDeviceToHost, copy_stream));
    OCU(c, cudaEventRecord(ob->landed, copy_stream));
    return NSB_OK;
}

// cs[0..n-1] = the slabs of one decomposition in rank order (n == 1: a plain context).
int nsb_out_wait(nsb_ctx* c0) {
    OutState* st = (OutState*)c0->out_state;
    if (!st) return NSB_OK;
    if (st->worker.joinable()) st->worker.join();
    st->running = false;
    if (st->status != NSB_OK) return nsb_fail(c0, st->status, "%s", st->err);
    return NSB_OK;
}

int nsb_out_run(nsb_ctx** cs, int n, double t, const char* prefix, int async) {
    nsb_ctx* c0 = cs[0];
    int r = nsb_out_wait(c0);                                  // one dump in flight per decomposition
    if (r != NSB_OK) return r;
    std::vector<OutSlab> slabs;
    for (int k = 0; k < n; k++) {
        if ((r = out_enqueue(cs[k], async != 0)) != NSB_OK) return r;
        slabs.push_back(out_rows(cs[k]));
    }
    OutState* st = (OutState*)c0->out_state;
    if (!st) { st = new OutState(); c0->out_state = st; }
    st->status = NSB_OK; st->err[0] = 0;
    const double a = c0->P.a, b = c0->P.b;
    const std::string pre(prefix);
    auto host_part = [=]() {
        for (const OutSlab& s : slabs) {
            cudaSetDevice(s.c->device);
            if (cudaEventSynchronize(((OutBuffers*)s.c->out_buffers)->landed) != cudaSuccess) {
                st->status = NSB_ERR_CUDA; snprintf(st->err, sizeof st->err, "output: the copy to the host failed");
                return;
            }
        }
        st->status = out_format_and_write(slabs, t, a, b, pre, st->err, sizeof st->err);
    };
    if (async) {
        st->running = true;
        st->worker = std::thread(host_part);
        return NSB_OK;
    }
    host_part();
    if (st->status != NSB_OK) return nsb_fail(c0, st->status, "%s", st->err);
    return NSB_OK;
}

// io.h:31 output(i_max, j_max, u, v, p, t, a, b, prefix) on the fields of a context
extern "C" int nsb_output(nsb_ctx* c, double t, const char* prefix) {
    if (!c || !prefix) return NSB_ERR_ARG;
    return nsb_out_run(&c, 1, t, prefix, 0);
}
extern "C" int nsb_output_async(nsb_ctx* c, double t, const char* prefix) {
    if (!c || !prefix) return NSB_ERR_ARG;
    return nsb_out_run(&c, 1, t, prefix, 1);
}
extern "C" int nsb_output_wait(nsb_ctx* c) {
    if (!c) return NSB_ERR_ARG;
    return nsb_out_wait(c);
}
