#include "pipeline.h"

#include "../../../../include/depthconv_b200.h"
#include "../../sens_loc/io/png.h"
#include "console.h"

#include <algorithm>
#include <atomic>
#include <chrono>
#include <condition_variable>
#include <cstdio>
#include <cstdlib>
#include <deque>
#include <functional>
#include <mutex>
#include <thread>
#include <sys/stat.h>

#include <sys/resource.h>
#include <sys/syscall.h>
#include <unistd.h>

namespace sens_loc::apps {
namespace {

namespace png = sens_loc::io::png;

void prof(const char* what, int a = -1) { profile_mark(what, a); }
}  // namespace

// DEPTH2X_PROFILE=1: stage timestamps on stderr (seconds since the first call)
void profile_mark(const char* what, int a) {
    static const bool on = std::getenv("DEPTH2X_PROFILE") != nullptr;
    if (!on)
        return;
    static const auto t0 = std::chrono::steady_clock::now();
    static std::mutex m;
    const double      t = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    std::lock_guard<std::mutex> l(m);
    std::fprintf(stderr, "[profile %8.4f s] %s %d\n", t, what, a);
}

namespace {

// ---- a persistent thread pool with a blocking parallel_for -------------------
class pool {
  public:
    explicit pool(int n) {
        for (int i = 0; i < std::max(1, n); ++i)
            _workers.emplace_back([this] { work(); });
    }
    ~pool() {
        {
            std::lock_guard<std::mutex> l(_m);
            _stop = true;
        }
        _cv.notify_all();
        for (auto& t : _workers)
            t.join();
    }
    /// fn(i, worker) for i in [0, n); returns when all calls are done.
    void parallel_for(int n, const std::function<void(int, int)>& fn) {
        if (n <= 0)
            return;
        std::unique_lock<std::mutex> l(_m);
        _fn = &fn, _n = n, _next = 0, _pending = n;
        ++_epoch;
        _cv.notify_all();
        _done.wait(l, [this] { return _pending == 0; });
        _fn = nullptr;
    }
    [[nodiscard]] int size() const { return int(_workers.size()); }

  private:
    void work() {
        // codec workers yield to the thread that initialises CUDA and feeds the GPU
        setpriority(PRIO_PROCESS, static_cast<id_t>(syscall(SYS_gettid)), 5);
        const int    me = _ids++;
        std::uint64_t seen = 0;
        std::unique_lock<std::mutex> l(_m);
        for (;;) {
            _cv.wait(l, [&] { return _stop || (_epoch != seen && _next < _n); });
            if (_stop)
                return;
            while (_fn != nullptr && _next < _n) {
                const int i = _next++;
                const auto* fn = _fn;
                l.unlock();
                (*fn)(i, me);
                l.lock();
                if (--_pending == 0)
                    _done.notify_all();
            }
            seen = _epoch;
        }
    }
    std::vector<std::thread>             _workers;
    std::mutex                           _m;
    std::condition_variable              _cv, _done;
    const std::function<void(int, int)>* _fn = nullptr;
    int                                  _n = 0, _next = 0, _pending = 0;
    std::uint64_t                        _epoch = 0;
    std::atomic<int>                     _ids{0};
    bool                                 _stop = false;
};

template <typename T>
class channel {
  public:
    void push(T v) {
        {
            std::lock_guard<std::mutex> l(_m);
            _q.push_back(std::move(v));
        }
        _cv.notify_one();
    }
    T pop() {
        std::unique_lock<std::mutex> l(_m);
        _cv.wait(l, [this] { return !_q.empty(); });
        T v = std::move(_q.front());
        _q.pop_front();
        return v;
    }

  private:
    std::mutex              _m;
    std::condition_variable _cv;
    std::deque<T>           _q;
};

// Batch buffers are ordinary page-aligned memory: dc_convert_host stages
// pageable memory through the context's own pinned ring (one memcpy per chunk on
// the otherwise idle GPU thread), whereas pinning 100-500 MB per buffer up front
// costs 0.1 s each on the reader's critical path, which dominates short jobs.
struct hostbuf {
    void*       p   = nullptr;
    std::size_t cap = 0;
    bool reserve(std::size_t bytes) {
        if (cap >= bytes)
            return true;
        std::free(p);
        p = nullptr;
        const std::size_t rounded = (bytes + 4095) & ~std::size_t(4095);
        if (posix_memalign(&p, 4096, rounded) != 0)
            p = nullptr;
        cap = p != nullptr ? rounded : 0;
        return p != nullptr;
    }
    hostbuf() = default;
    hostbuf(const hostbuf&)            = delete;
    hostbuf& operator=(const hostbuf&) = delete;
    ~hostbuf() { std::free(p); }
};

struct batch {
    hostbuf          in, out[DC_MAX_OUTPUTS];
    std::vector<int> idx;  // index of the frame in slot k
    int              w = 0, h = 0;
    bool             last = false, gpu_ok = true;
};

struct shared_state {
    std::mutex       m;
    std::vector<int> failed;
    void fail(int idx) {
        std::lock_guard<std::mutex> l(m);
        failed.push_back(idx);
        // parallel_processing.h:46-54
        std::cerr << console::err() << "Could not process index \"" << console::bold() << idx
                  << "\"" << console::reset() << "!" << std::endl;
    }
};

constexpr std::size_t kMaxFramePixels = std::size_t(1) << 28;  // 16384 x 16384

struct pending_file {
    int                        idx = 0;
    std::vector<unsigned char> bytes;
    png::info                  hdr;
    bool                       ok = false;
};

// One GPU: indices [first, last] through reader -> gpu -> writer.
void run_device(const convert_spec& spec, int device, int first, int last, int batch_size,
                int threads, shared_state& st) {
    // The CUDA context (cuInit + context: 0.4 - 1.5 s of a fresh process, ten times
    // the pipeline time of a 100-frame job) is created by the GPU thread while the
    // reader already loads and decodes the first batches: the start-up overlaps the
    // decode instead of preceding it. DEPTH2X_SERIAL_INIT=1 restores the old order
    // (context first), which measured faster only when the codec threads' first
    // touches of large buffers fight cuInit for the address-space lock.
    static const bool serial_init = std::getenv("DEPTH2X_SERIAL_INIT") != nullptr;
    dc_ctx*           ctx         = nullptr;
    bool              have_ctx    = false;
    auto create_context = [&] {
        prof("dc_create begin", device);
        have_ctx = dc_create(device, &ctx) == DC_OK;
        prof("dc_create end", device);
        if (!have_ctx) {
            std::lock_guard<std::mutex> l(st.m);
            std::cerr << console::err() << "GPU " << device << ": " << dc_last_error(nullptr) << "\n";
        }
    };
    if (serial_init)
        create_context();
    int n_out = 0;
    for (const auto& o : spec.output)
        n_out += !o.empty();

    constexpr int   kBuffers = 4;
    batch           buffers[kBuffers];
    channel<batch*> free_q, gpu_q, write_q;
    for (auto& b : buffers)
        free_q.push(&b);
    // Both pools get every thread: decode and encode work are unequal (bearing
    // writes four files per frame read) and the kernel time-slices what is runnable.
    pool read_pool(std::max(1, threads)), write_pool(std::max(1, threads));
    const int total = last - first + 1;

    // ---- reader ------------------------------------------------------------
    std::thread reader([&] {
        std::vector<std::vector<unsigned char>> scratch;
        scratch.resize(std::size_t(read_pool.size()));
        int next = first;
        while (next <= last) {
            // frames per batch: at most 256 MB per buffer and 64 frames, at least
            // one frame per thread, about six batches per job so that decoding, the
            // GPU and encoding overlap even on short jobs, a multiple of the thread count
            int want = batch_size;
            if (want <= 0) {
                const std::size_t px = spec.camera.model != 0
                                           ? std::size_t(spec.camera.w) * spec.camera.h
                                           : std::size_t(1920) * 1080;
                const int cap = int(std::clamp<std::size_t>((std::size_t(256) << 20) / (px * 2 * (1 + n_out)), 1, 64));
                want = std::clamp(std::max((total + 5) / 6, threads), 1, cap);
                // whole rounds of the codec pool: 17 frames on 16 threads decode and
                // encode in two rounds, the second with one busy thread (measured on
                // BASELINE config 1: 100 frames -> batches of 17)
                if (want > threads)
                    want -= want % threads;
            }
            const int n = std::min(want, last - next + 1);
            std::vector<pending_file> files{};
            files.resize(std::size_t(n));
            // phase A: read the files, look at the headers (io/image.h:21-31: the
            // image must exist and be 16 bit, one channel)
            read_pool.parallel_for(n, [&](int k, int) {
                pending_file& f = files[std::size_t(k)];
                f.idx           = next + k;
                std::string path;
                if (!format_index(spec.input, f.idx, path))
                    return;
                if (png::read_file(path, f.bytes, f.hdr) != png::status::ok)
                    return;
                // (the header is untrusted: a frame beyond 2^28 pixels is refused before
                // any buffer is sized from it)
                f.ok = f.hdr.color_type == 0 && f.hdr.bit_depth == 16 && f.hdr.interlace == 0 &&
                       f.hdr.w > 0 && f.hdr.h > 0 &&
                       std::size_t(f.hdr.w) * std::size_t(f.hdr.h) <= kMaxFramePixels;
            });
            // frame size of this batch: the intrinsic's (batch_converter.h:152-154),
            // for `scale` the size of the first readable frame; a frame of another
            // size ends the batch (scale) or fails (camera)
            int bw = spec.camera.w, bh = spec.camera.h, use = n;
            if (spec.camera.model == 0) {
                bw = bh = 0;
                for (int k = 0; k < n; ++k) {
                    const pending_file& f = files[std::size_t(k)];
                    if (!f.ok)
                        continue;
                    if (bw == 0)
                        bw = f.hdr.w, bh = f.hdr.h;
                    else if (f.hdr.w != bw || f.hdr.h != bh) {
                        use = k;
                        break;
                    }
                }
            }
            prof("files read", n);
            batch* b = free_q.pop();
            b->idx.clear();
            b->w = bw, b->h = bh;
            b->last   = next + use > last;
            b->gpu_ok = true;
            std::vector<int> slot_of;
            slot_of.assign(std::size_t(use), -1);
            for (int k = 0; k < use; ++k) {
                pending_file& f = files[std::size_t(k)];
                if (f.ok && (f.hdr.w != bw || f.hdr.h != bh))
                    f.ok = false;
                if (f.ok) {
                    slot_of[std::size_t(k)] = int(b->idx.size());
                    b->idx.push_back(f.idx);
                } else {
                    st.fail(f.idx);
                }
            }
            const std::size_t fpx = std::size_t(bw) * std::size_t(bh);
            bool              mem = b->in.reserve(std::max<std::size_t>(1, b->idx.size() * fpx * 2));
            for (int s = 0, o = 0; s < DC_MAX_OUTPUTS; ++s)
                if (!spec.output[s].empty())
                    mem = mem && b->out[o++].reserve(std::max<std::size_t>(1, b->idx.size() * fpx * 2));
            std::vector<char> bad;
            bad.assign(std::size_t(use), 0);
            prof("pinned ready", use);
            if (mem) {
                // phase B: inflate + unfilter straight into the pinned batch buffer
                read_pool.parallel_for(use, [&](int k, int worker) {
                    const int s = slot_of[std::size_t(k)];
                    if (s < 0)
                        return;
                    pending_file& f   = files[std::size_t(k)];
                    auto*         dst = static_cast<unsigned short*>(b->in.p) + std::size_t(s) * fpx;
                    if (png::decode_grey(f.bytes, f.hdr, dst, scratch[std::size_t(worker)]) != png::status::ok)
                        bad[std::size_t(k)] = 1;
                });
            }
            // a corrupt stream fails its index; its slot still travels (zeros are
            // valid input) but is not written
            for (int k = 0; k < use; ++k)
                if (slot_of[std::size_t(k)] >= 0 && (bad[std::size_t(k)] || !mem)) {
                    st.fail(files[std::size_t(k)].idx);
                    b->idx[std::size_t(slot_of[std::size_t(k)])] = -1 - files[std::size_t(k)].idx;
                }
            next += use;
            prof("decoded", use);
            gpu_q.push(b);
        }
    });

    // ---- gpu -----------------------------------------------------------------
    std::thread gpu([&] {
        if (!serial_init)
            create_context();
        int cam_w = -1, cam_h = -1;
        for (;;) {
            batch* b = gpu_q.pop();
            if (!have_ctx) {
                b->gpu_ok = false;  // every index of the batch is reported by the writer
            } else if (!b->idx.empty()) {
                int rc = DC_OK;
                if (spec.camera.model == 0) {
                    if (b->w != cam_w || b->h != cam_h)
                        rc = dc_set_frame_size(ctx, b->w, b->h);
                } else if (cam_w < 0) {
                    rc = spec.camera.model == 1
                             ? dc_set_pinhole(ctx, spec.camera.w, spec.camera.h, spec.camera.p[0],
                                              spec.camera.p[1], spec.camera.p[2], spec.camera.p[3])
                             : dc_set_equirectangular(ctx, spec.camera.w, spec.camera.h,
                                                      spec.camera.p[0], spec.camera.p[1]);
                }
                if (rc == DC_OK)
                    cam_w = b->w, cam_h = b->h;  // (a failed set is retried by the next batch)
                if (rc == DC_OK) {
                    dc_job j{};
                    j.feature  = spec.feature;
                    j.in_kind  = spec.in_kind;
                    j.out_kind = DC_OUT_U16;
                    j.n_frames = int(b->idx.size());
                    j.in       = b->in.p;
                    for (int s = 0, o = 0; s < DC_MAX_OUTPUTS; ++s)
                        if (!spec.output[s].empty())
                            j.out[s] = b->out[o++].p;
                    j.clamp_lo = spec.clamp_lo, j.clamp_hi = spec.clamp_hi;
                    j.scale = spec.scale, j.offset = spec.offset;
                    j.median_ksize = spec.median;
                    prof("camera set", j.n_frames);
                    rc = dc_convert_host(ctx, &j);
                    prof("converted", j.n_frames);
                }
                if (rc != DC_OK) {
                    std::lock_guard<std::mutex> l(st.m);
                    std::cerr << console::err() << "GPU " << device << ": " << dc_last_error(ctx) << "\n";
                    b->gpu_ok = false;
                }
            }
            const bool last_batch = b->last;
            write_q.push(b);
            if (last_batch)
                return;
        }
    });

    // ---- writer --------------------------------------------------------------
    std::thread writer([&] {
        std::vector<std::vector<unsigned char>> scratch, encoded;
        scratch.resize(std::size_t(write_pool.size()));
        encoded.resize(std::size_t(write_pool.size()));
        for (;;) {
            batch*            b   = write_q.pop();
            const int         n   = int(b->idx.size());
            const std::size_t fpx = std::size_t(b->w) * std::size_t(b->h);
            std::vector<char> bad;
            bad.assign(std::size_t(n), 0);
            int               slots[DC_MAX_OUTPUTS], n_slots = 0;
            for (int s = 0; s < DC_MAX_OUTPUTS; ++s)
                if (!spec.output[s].empty())
                    slots[n_slots++] = s;
            if (b->gpu_ok) {
                write_pool.parallel_for(n * n_slots, [&](int t, int worker) {
                    const int k = t / n_slots, o = t % n_slots;
                    if (b->idx[std::size_t(k)] < 0)
                        return;
                    std::string path;
                    const auto* src = static_cast<const unsigned short*>(b->out[o].p) + std::size_t(k) * fpx;
                    if (!format_index(spec.output[slots[o]], b->idx[std::size_t(k)], path) ||
                        png::encode_grey(src, b->w, b->h, 16, encoded[std::size_t(worker)],
                                         scratch[std::size_t(worker)]) != png::status::ok ||
                        png::write_file(path, encoded[std::size_t(worker)]) != png::status::ok)
                        bad[std::size_t(k)] = 1;
                });
            }
            for (int k = 0; k < n; ++k)
                if (b->idx[std::size_t(k)] >= 0 && (bad[std::size_t(k)] || !b->gpu_ok))
                    st.fail(b->idx[std::size_t(k)]);
            prof("encoded+written", n);
            const bool last_batch = b->last;
            free_q.push(b);
            if (last_batch)
                return;
        }
    });

    reader.join();
    gpu.join();
    writer.join();
    prof("pipeline drained", device);
    if (have_ctx)
        dc_destroy(ctx);
    prof("context destroyed", device);
}

// GPUs this process can use, without touching the CUDA driver: the entries of a
// CUDA_VISIBLE_DEVICES list the caller set, else the /dev/nvidia<N> device nodes.
// 0 = cannot tell.
int visible_gpus_without_cuda(bool caller_chose) {
    if (caller_chose) {
        const std::string v = std::getenv("CUDA_VISIBLE_DEVICES");
        if (v.empty())
            return 0;
        int n = 1;
        for (char c : v)
            n += c == ',';
        return n;
    }
    int n = 0;
    for (int i = 0; i < 64; ++i) {
        struct stat sb;
        if (::stat(("/dev/nvidia" + std::to_string(i)).c_str(), &sb) == 0)
            ++n;
    }
    return n;
}

}  // namespace

run_stats run_batch(const convert_spec& spec, int start, int end, int devices, int batch_size,
                    int threads) {
    if (start > end)
        std::swap(start, end);  // parallel_processing.h:31-34
    const int total = end - start + 1;
    if (threads <= 0)
        threads = int(std::max(2u, std::thread::hardware_concurrency()));
    devices = std::max(1, std::min(devices, total));
    // cuInit time grows with the number of visible GPUs: expose only the ones used
    // (unless the caller already chose).
    const bool caller_chose = std::getenv("CUDA_VISIBLE_DEVICES") != nullptr;
    if (!caller_chose) {
        std::string vis;
        for (int g = 0; g < devices; ++g)
            vis += (g ? "," : "") + std::to_string(g);
        setenv("CUDA_VISIBLE_DEVICES", vis.c_str(), 0);
    }
    if (devices > 1) {
        // How many GPUs are there? Asked WITHOUT initialising CUDA (dc_device_count()
        // costs the whole cuInit, 1.5-2.5 s, before the first file could be read:
        // measured on a 2-GPU box, the first frame was decoded 2.3 s after start
        // instead of 0.03 s): the caller's CUDA_VISIBLE_DEVICES list, else the
        // /dev/nvidiaN nodes; only if neither tells, the CUDA query.
        int visible = visible_gpus_without_cuda(caller_chose);
        if (visible <= 0)
            visible = dc_device_count();
        if (visible > 0)
            devices = std::min(devices, visible);
    }

    shared_state             st;
    const auto               t0 = std::chrono::steady_clock::now();
    std::vector<std::thread> workers;
    // contiguous blocks (sequential file access per GPU), the first total % devices
    // of them one index longer: every GPU gets work (sharding.py contiguous_blocks)
    const int base = total / devices, extra = total % devices;
    for (int g = 0, first = start; g < devices; ++g) {
        const int n = base + (g < extra ? 1 : 0), last = first + n - 1;
        workers.emplace_back([&, g, first, last] {
            run_device(spec, g, first, last, batch_size, std::max(2, threads / devices), st);
        });
        first += n;
    }
    for (auto& w : workers)
        w.join();
    run_stats r;
    r.seconds   = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    r.failed    = int(st.failed.size());
    r.processed = total - r.failed;
    return r;
}

}  // namespace sens_loc::apps
