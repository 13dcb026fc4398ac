// Process-wide pinned-host cache of layer files + safetensors header parsing.
//
// Stands in for the reference's file handling: `safetensors_load(f.read())` per layer per call
// (reference utils.py:126-127) and, in data-parallel mode, DeviceManager's "read each layer to RAM
// once and hand the same state dict to every GPU thread" (utils.py:35-41,67-75).  Here each layer
// file is read once per process into page-locked memory (parallel pread) and every context copies
// from that image, so H2D copies are asynchronous and run at link rate.
#include "host_cache.h"

#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>
#include <cerrno>
#include <climits>
#include <cstdlib>

#include <algorithm>
#include <atomic>
#include <functional>
#include <cstring>
#include <mutex>
#include <thread>
#include <unordered_map>

#include <cuda_fp16.h>
#include <cuda_runtime.h>

#include "common.cuh"

namespace fls {

namespace {

// ---- a JSON reader just big enough for a safetensors header --------------------------------
struct Json {
    const char* p;
    const char* end;
    bool ok = true;

    void ws() {
        while (p < end && (*p == ' ' || *p == '\n' || *p == '\t' || *p == '\r')) ++p;
    }
    bool eat(char c) {
        ws();
        if (p < end && *p == c) {
            ++p;
            return true;
        }
        return false;
    }
    std::string str() {
        ws();
        std::string s;
        if (p >= end || *p != '"') {
            ok = false;
            return s;
        }
        ++p;
        while (p < end && *p != '"') {
            if (*p == '\\' && p + 1 < end) {
                ++p;
                switch (*p) {
                    case 'n': s += '\n'; break;
                    case 't': s += '\t'; break;
                    case 'u': s += '?'; p += std::min<ptrdiff_t>(4, end - p - 1); break;
                    default: s += *p;
                }
                ++p;
            } else {
                s += *p++;
            }
        }
        if (p >= end) ok = false;
        else ++p;
        return s;
    }
    long long num() {
        ws();
        char* e = nullptr;
        long long v = strtoll(p, &e, 10);
        if (e == p) ok = false;
        p = e;
        return v;
    }
    void skip() {  // skip any value
        ws();
        if (p >= end) {
            ok = false;
            return;
        }
        if (*p == '"') {
            str();
        } else if (*p == '{' || *p == '[') {
            const char close = *p == '{' ? '}' : ']';
            const bool obj = *p == '{';
            ++p;
            if (eat(close)) return;
            do {
                if (obj) {
                    str();
                    if (!eat(':')) ok = false;
                }
                skip();
            } while (ok && eat(','));
            if (!eat(close)) ok = false;
        } else {
            while (p < end && *p != ',' && *p != '}' && *p != ']') ++p;
        }
    }
};

struct CacheEntry {
    std::shared_ptr<HostLayerImage> img;
    uint64_t last_use = 0;
};

std::mutex g_mu;
std::unordered_map<std::string, CacheEntry> g_cache;
std::atomic<uint64_t> g_clock{0};
std::atomic<uint64_t> g_file_bytes{0};
std::atomic<uint64_t> g_pinned_bytes{0};
std::atomic<uint64_t> g_mapped_bytes{0};
std::atomic<size_t> g_limit{0};
std::atomic<int> g_image_mode{-1};  // -1: not set yet -> FLS_WEIGHT_IMAGE (auto | copy | map)

int image_mode() {
    int m = g_image_mode.load();
    if (m < 0) {
        const char* e = getenv("FLS_WEIGHT_IMAGE");
        m = !e ? 0 : (strcmp(e, "copy") == 0 ? 1 : (strcmp(e, "map") == 0 ? 2 : 0));
        g_image_mode.store(m);
    }
    return m;
}

// Touch every page of [p, p+n) from a few threads so that the file is read (or the tmpfs pages are found) in
// parallel before cudaHostRegister walks the range on one thread.
void prefault(const char* p, size_t n) {
    const size_t chunk = size_t(64) << 20;
    const size_t n_chunks = (n + chunk - 1) / chunk;
    const unsigned hw = std::max(1u, std::thread::hardware_concurrency());
    const size_t n_threads = std::min<size_t>(std::min<size_t>(8, hw), std::max<size_t>(1, n_chunks));
    std::atomic<size_t> next{0};
    std::atomic<unsigned> sink{0};
    auto work = [&]() {
        unsigned acc = 0;
        for (;;) {
            const size_t c = next.fetch_add(1);
            if (c >= n_chunks) break;
            const size_t lo = c * chunk, hi = std::min(n, lo + chunk);
            for (size_t o = lo; o < hi; o += 4096) acc += static_cast<unsigned char>(p[o]);
        }
        sink.fetch_add(acc);
    };
    std::vector<std::thread> th;
    for (size_t i = 1; i < n_threads; ++i) th.emplace_back(work);
    work();
    for (auto& t : th) t.join();
}

int parallel_pread(int fd, char* dst, size_t off, size_t n);

std::mutex g_stage_mu;
std::string g_stage_dir;   // empty = no staging

// Copy of a disk-resident layer file in the staging directory (tmpfs), made once per BOX: the first process to get
// there creates `<name>.lock`, copies the file (past the page cache where the file system allows) into `<name>.tmp`
// and renames it to `<name>`; the others wait for `<name>` to appear.  The name carries the source's size and
// modification time, so a changed source file is staged afresh.  Returns false (and says why) when staging is off, the
// directory cannot take the file, or the copying process does not finish in time - the caller then copies privately.
bool stage_file(const std::string& src, size_t file_size, std::string* staged, std::string* why) {
    std::string dir;
    {
        std::lock_guard<std::mutex> lk(g_stage_mu);
        dir = g_stage_dir;
    }
    if (dir.empty()) {
        *why = "no staging directory";
        return false;
    }
    struct stat st;
    if (stat(src.c_str(), &st) != 0) {
        *why = std::string("stat: ") + strerror(errno);
        return false;
    }
    const size_t h = std::hash<std::string>()(src);
    char name[160];
    const char* base = strrchr(src.c_str(), '/');
    snprintf(name, sizeof(name), "%016zx-%zu-%lld-%.80s", h, file_size, static_cast<long long>(st.st_mtime),
             base ? base + 1 : src.c_str());
    const std::string final_path = dir + "/" + name, lock_path = final_path + ".lock", tmp_path = final_path + ".tmp";
    auto exists = [](const std::string& p) { struct stat s2; return stat(p.c_str(), &s2) == 0; };
    if (exists(final_path)) {
        *staged = final_path;
        return true;
    }
    const int lock_fd = open(lock_path.c_str(), O_CREAT | O_EXCL | O_WRONLY, 0600);
    if (lock_fd < 0) {
        if (errno != EEXIST) {
            *why = std::string("staging directory: ") + strerror(errno);
            return false;
        }
        for (int i = 0; i < 6000; ++i) {   // another process is copying: up to 10 minutes
            if (exists(final_path)) {
                *staged = final_path;
                return true;
            }
            if (!exists(lock_path)) break;   // it gave up
            usleep(100000);
        }
        *why = "the process staging this file did not finish";
        return false;
    }
    close(lock_fd);
    bool ok = false;
    const int out_fd = open(tmp_path.c_str(), O_CREAT | O_TRUNC | O_RDWR, 0600);
    if (out_fd >= 0 && ftruncate(out_fd, static_cast<off_t>(file_size)) == 0) {
        const size_t page = static_cast<size_t>(sysconf(_SC_PAGESIZE));
        const size_t len = (file_size + page - 1) / page * page;
        void* m = mmap(nullptr, len, PROT_READ | PROT_WRITE, MAP_SHARED, out_fd, 0);
        if (m != MAP_FAILED) {
            const int in_fd = open(src.c_str(), O_RDONLY);
            if (in_fd >= 0) {
                ok = parallel_pread(in_fd, static_cast<char*>(m), 0, file_size) == 0;
                posix_fadvise(in_fd, 0, 0, POSIX_FADV_DONTNEED);   // tmpfs now holds the bytes: not twice in RAM
                close(in_fd);
                if (ok) g_file_bytes.fetch_add(file_size);
            }
            munmap(m, len);
        }
    }
    if (out_fd >= 0) close(out_fd);
    if (ok) ok = rename(tmp_path.c_str(), final_path.c_str()) == 0;
    if (!ok) {
        *why = std::string("copy into the staging directory failed: ") + strerror(errno);
        unlink(tmp_path.c_str());
    }
    unlink(lock_path.c_str());
    if (ok) *staged = final_path;
    return ok;
}

// Tries to make `img` the mapped form of its file.  Returns true on success; false (with CUDA's error state
// cleared) when the mapping or the registration is refused - the caller then falls back to the copied form.
bool try_map_image(HostLayerImage* img, const std::string& map_path, size_t data_start, size_t file_size,
                   std::string* why) {
    // What this driver accepts (tools/hostreg_probe.cu, B200 box, driver 580): the mapping must be WRITABLE even for
    // a read-only registration, and file-backed MAP_SHARED pages are accepted on tmpfs (/dev/shm) only - on ext4
    // every MAP_SHARED variant is refused and MAP_PRIVATE works only by copying each page (no sharing, no saving).
    // So: open read-write, map shared, never write.  Anything else falls back to the copying path.
    const int fd = open(map_path.c_str(), O_RDWR);
    if (fd < 0) {
        *why = std::string("open for mapping: ") + strerror(errno);
        return false;
    }
    const size_t page = static_cast<size_t>(sysconf(_SC_PAGESIZE));
    const size_t len = (file_size + page - 1) / page * page;
    void* m = mmap(nullptr, len, PROT_READ | PROT_WRITE, MAP_SHARED, fd, 0);
    close(fd);
    if (m == MAP_FAILED) {
        *why = std::string("mmap: ") + strerror(errno);
        return false;
    }
    prefault(static_cast<const char*>(m), file_size);
    cudaError_t e = cudaHostRegister(m, len, cudaHostRegisterPortable | cudaHostRegisterReadOnly);
    if (e != cudaSuccess) {
        cudaGetLastError();
        *why = std::string("cudaHostRegister: ") + cudaGetErrorString(e) +
               " (file pages can be locked in place on tmpfs, e.g. /dev/shm, only)";
        munmap(m, len);
        return false;
    }
    img->map_base = m;
    img->map_len = len;
    img->pinned = static_cast<char*>(m) + data_start;
    img->pinned_bytes = len;
    for (auto& t : img->tensors) t.host_off = t.begin;
    g_pinned_bytes.fetch_add(len);
    g_mapped_bytes.fetch_add(len);
    if (map_path == img->path) g_file_bytes.fetch_add(file_size - data_start);   // (a staged copy was counted when made)
    return true;
}

int parallel_pread(int fd, char* dst, size_t off, size_t n) {
    const size_t chunk = size_t(32) << 20;
    const size_t n_chunks = (n + chunk - 1) / chunk;
    const unsigned hw = std::max(1u, std::thread::hardware_concurrency());
    const size_t n_threads = std::min<size_t>(std::min<size_t>(8, hw), std::max<size_t>(1, n_chunks));
    std::atomic<size_t> next{0};
    std::atomic<int> err{0};
    auto work = [&]() {
        for (;;) {
            const size_t c = next.fetch_add(1);
            if (c >= n_chunks || err.load()) return;
            size_t o = c * chunk, left = std::min(chunk, n - o);
            while (left) {
                const ssize_t r = pread(fd, dst + o, left, static_cast<off_t>(off + o));
                if (r <= 0) {
                    err.store(1);
                    return;
                }
                o += static_cast<size_t>(r);
                left -= static_cast<size_t>(r);
            }
        }
    };
    std::vector<std::thread> th;
    for (size_t i = 1; i < n_threads; ++i) th.emplace_back(work);
    work();
    for (auto& t : th) t.join();
    return err.load() ? -1 : 0;
}

void convert_to_f16(const char* src, const std::string& dtype, size_t n, __half* dst) {
    if (dtype == "F32") {
        const float* s = reinterpret_cast<const float*>(src);
        for (size_t i = 0; i < n; ++i) dst[i] = __float2half_rn(s[i]);
    } else {  // BF16
        const uint16_t* s = reinterpret_cast<const uint16_t*>(src);
        for (size_t i = 0; i < n; ++i) {
            const uint32_t bits = static_cast<uint32_t>(s[i]) << 16;
            float f;
            memcpy(&f, &bits, 4);
            dst[i] = __float2half_rn(f);
        }
    }
}

}  // namespace

HostLayerImage::~HostLayerImage() {
    if (map_base) {
        cudaHostUnregister(map_base);
        munmap(map_base, map_len);
        g_pinned_bytes.fetch_sub(pinned_bytes);
        g_mapped_bytes.fetch_sub(pinned_bytes);
    } else if (pinned) {
        cudaFreeHost(pinned);
        g_pinned_bytes.fetch_sub(pinned_bytes);
    }
}

int parse_safetensors_header(const std::string& path, std::vector<TensorEntry>* out, size_t* data_start,
                             size_t* file_size) {
    const int fd = open(path.c_str(), O_RDONLY);
    if (fd < 0) return fail(kErrIo, "cannot open %s: %s", path.c_str(), strerror(errno));
    struct stat st;
    fstat(fd, &st);
    uint64_t hlen = 0;
    if (pread(fd, &hlen, 8, 0) != 8 || hlen == 0 || hlen > (100u << 20) ||
        8 + hlen > static_cast<uint64_t>(st.st_size)) {
        close(fd);
        return fail(kErrIo, "%s: not a safetensors file (bad header length)", path.c_str());
    }
    std::string hdr(hlen, '\0');
    if (pread(fd, hdr.data(), hlen, 8) != static_cast<ssize_t>(hlen)) {
        close(fd);
        return fail(kErrIo, "%s: short read of the header", path.c_str());
    }
    close(fd);
    *data_start = 8 + hlen;
    *file_size = static_cast<size_t>(st.st_size);

    Json j{hdr.data(), hdr.data() + hdr.size()};
    if (!j.eat('{')) return fail(kErrIo, "%s: header is not a JSON object", path.c_str());
    if (!j.eat('}')) {
        do {
            const std::string name = j.str();
            if (!j.eat(':')) j.ok = false;
            if (name == "__metadata__") {
                j.skip();
                continue;
            }
            TensorEntry t;
            t.name = name;
            if (!j.eat('{')) j.ok = false;
            do {
                const std::string key = j.str();
                if (!j.eat(':')) j.ok = false;
                if (key == "dtype") {
                    t.dtype = j.str();
                } else if (key == "shape") {
                    if (!j.eat('[')) j.ok = false;
                    if (!j.eat(']')) {
                        do t.shape.push_back(j.num());
                        while (j.ok && j.eat(','));
                        if (!j.eat(']')) j.ok = false;
                    }
                } else if (key == "data_offsets") {
                    if (!j.eat('[')) j.ok = false;
                    t.begin = static_cast<size_t>(j.num());
                    if (!j.eat(',')) j.ok = false;
                    t.end = static_cast<size_t>(j.num());
                    if (!j.eat(']')) j.ok = false;
                } else {
                    j.skip();
                }
            } while (j.ok && j.eat(','));
            if (!j.eat('}')) j.ok = false;
            if (!j.ok) break;
            size_t n = 1;
            for (long long d : t.shape) n *= static_cast<size_t>(d);
            t.numel = n;
            if (t.end < t.begin || *data_start + t.end > *file_size)
                return fail(kErrIo, "%s: tensor %s data_offsets outside the file", path.c_str(), name.c_str());
            out->push_back(std::move(t));
        } while (j.ok && j.eat(','));
        if (j.ok && !j.eat('}')) j.ok = false;
    }
    if (!j.ok) return fail(kErrIo, "%s: malformed safetensors header", path.c_str());
    return kOk;
}

namespace {
bool is_pack_path(const std::string& path) {
    static const char kExt[] = ".flspack";
    return path.size() > sizeof(kExt) - 1 && path.compare(path.size() - (sizeof(kExt) - 1), sizeof(kExt) - 1, kExt) == 0;
}
}  // namespace

// Packed layer file (prepare_weights.py --packed): 4096-byte header {"FLSPACK1", u32 kind, u32 n, u64 data_offset,
// u64 data_bytes, n x {char name[96]; u64 rows, cols, offset, nbytes}}, then fp16 tensors, each 256-byte aligned, in
// HBM slot order; data_bytes is a whole number of 4096-byte pages.
int parse_pack_header(const std::string& path, std::vector<TensorEntry>* out, size_t* data_start, size_t* file_size) {
    const int fd = open(path.c_str(), O_RDONLY);
    if (fd < 0) return fail(kErrIo, "cannot open %s: %s", path.c_str(), strerror(errno));
    struct stat st;
    fstat(fd, &st);
    unsigned char head[4096];
    const ssize_t got = pread(fd, head, sizeof(head), 0);
    close(fd);
    uint32_t n = 0;
    uint64_t data_off = 0, data_bytes = 0;
    if (got == static_cast<ssize_t>(sizeof(head))) {
        memcpy(&n, head + 12, 4);
        memcpy(&data_off, head + 16, 8);
        memcpy(&data_bytes, head + 24, 8);
    }
    if (got != static_cast<ssize_t>(sizeof(head)) || memcmp(head, "FLSPACK1", 8) != 0 || data_off != 4096 || n == 0 ||
        n > 31 || data_off + data_bytes != static_cast<uint64_t>(st.st_size) || data_bytes % 4096)
        return fail(kErrIo, "%s: not a packed layer file (bad header)", path.c_str());
    for (uint32_t i = 0; i < n; ++i) {
        const unsigned char* e = head + 32 + i * 128;
        uint64_t rows, cols, off, nb;
        memcpy(&rows, e + 96, 8);
        memcpy(&cols, e + 104, 8);
        memcpy(&off, e + 112, 8);
        memcpy(&nb, e + 120, 8);
        TensorEntry t;
        t.name.assign(reinterpret_cast<const char*>(e), strnlen(reinterpret_cast<const char*>(e), 96));
        t.dtype = "F16";
        t.shape.push_back(static_cast<long long>(rows));
        if (cols) t.shape.push_back(static_cast<long long>(cols));
        t.numel = static_cast<size_t>(rows) * (cols ? cols : 1);
        t.begin = off;
        t.end = off + nb;
        if (nb != t.numel * 2 || off % 256 || t.end > data_bytes)
            return fail(kErrIo, "%s: tensor %s has an inconsistent entry in the packed header", path.c_str(),
                        t.name.c_str());
        out->push_back(std::move(t));
    }
    *data_start = data_off;
    *file_size = static_cast<size_t>(st.st_size);
    return kOk;
}

int parse_layer_header(const std::string& path, std::vector<TensorEntry>* out, size_t* data_start, size_t* file_size) {
    return is_pack_path(path) ? parse_pack_header(path, out, data_start, file_size)
                              : parse_safetensors_header(path, out, data_start, file_size);
}

int acquire_host_image(const std::string& path_in, std::shared_ptr<HostLayerImage>* out) {
    // key by the resolved path: symlinked layer files share one pinned image
    std::string path = path_in;
    if (char* rp = realpath(path_in.c_str(), nullptr)) {
        path = rp;
        free(rp);
    }
    {
        std::lock_guard<std::mutex> lk(g_mu);
        auto it = g_cache.find(path);
        if (it != g_cache.end()) {
            it->second.last_use = ++g_clock;
            *out = it->second.img;
            return kOk;
        }
    }
    // Load outside the lock on a private image, publish afterwards (a racing duplicate is dropped).
    auto img = std::make_shared<HostLayerImage>();
    img->path = path;
    size_t data_start = 0, file_size = 0;
    FLS_TRY(parse_layer_header(path, &img->tensors, &data_start, &file_size));
    const bool packed = is_pack_path(path);
    size_t total = 0;
    bool all_f16 = true;
    for (auto& t : img->tensors) {
        if (t.dtype != "F16" && t.dtype != "BF16" && t.dtype != "F32")
            return fail(kErrInvalid, "%s: tensor %s has dtype %s; only F16/BF16/F32 weights are supported "
                        "(int8 needs fp16 statistics, as in the reference)", path.c_str(), t.name.c_str(),
                        t.dtype.c_str());
        all_f16 = all_f16 && t.dtype == "F16" && t.end - t.begin == t.numel * 2 && (data_start + t.begin) % 2 == 0;
        t.host_off = packed ? t.begin : total;     // packed: the image IS the file's data section
        total += (t.numel * 2 + 255) & ~size_t(255);
    }
    if (packed) total = file_size - data_start;
    const int mode = image_mode();
    bool mapped = false;
    if (mode != 1) {
        std::string why = "the file holds BF16/F32 tensors, which are converted on load";
        mapped = all_f16 && try_map_image(img.get(), path, data_start, file_size, &why);
        if (!mapped && all_f16) {
            // a file on a disk file system: one shared copy per box in the staging directory, if there is one
            std::string staged, why_stage;
            if (stage_file(path, file_size, &staged, &why_stage))
                mapped = try_map_image(img.get(), staged, data_start, file_size, &why);
            else if (why_stage != "no staging directory")
                why += "; staging: " + why_stage;
        }
        if (!mapped && mode == 2)
            return fail(kErrIo, "%s: cannot lock the file's pages in place (%s)", path.c_str(), why.c_str());
    }
    if (mapped) {
        std::lock_guard<std::mutex> lk(g_mu);
        auto& slot = g_cache[path];
        if (!slot.img) slot.img = img;
        slot.last_use = ++g_clock;
        *out = slot.img;
    }
    if (mapped) {
        img.reset();
        if (const size_t lim = g_limit.load()) trim_host_cache(lim);
        return kOk;
    }
    cudaError_t e = cudaHostAlloc(reinterpret_cast<void**>(&img->pinned), std::max<size_t>(total, 256),
                                  cudaHostAllocPortable);
    if (e != cudaSuccess)
        return fail(kErrCuda, "cudaHostAlloc(%zu bytes) for %s failed: %s", total, path.c_str(), cudaGetErrorString(e));
    img->pinned_bytes = total;
    g_pinned_bytes.fetch_add(total);

    if (packed) {
        // one page-aligned range straight into the page-locked buffer, past the page cache where the file system
        // allows it (O_DIRECT; tmpfs does not)
        int fd = open(path.c_str(), O_RDONLY | O_DIRECT);
        const bool direct = fd >= 0;
        if (!direct) fd = open(path.c_str(), O_RDONLY);
        if (fd < 0) return fail(kErrIo, "cannot open %s: %s", path.c_str(), strerror(errno));
        const int prc = parallel_pread(fd, img->pinned, data_start, total);
        if (!direct) posix_fadvise(fd, 0, 0, POSIX_FADV_DONTNEED);
        close(fd);
        if (prc) return fail(kErrIo, "%s: short read of the packed data", path.c_str());
        g_file_bytes.fetch_add(total);
        std::lock_guard<std::mutex> lk(g_mu);
        auto& slot = g_cache[path];
        if (!slot.img) slot.img = img;
        slot.last_use = ++g_clock;
        *out = slot.img;
    }
    if (packed) {
        img.reset();
        if (const size_t lim = g_limit.load()) trim_host_cache(lim);
        return kOk;
    }
    const int fd = open(path.c_str(), O_RDONLY);
    if (fd < 0) return fail(kErrIo, "cannot open %s: %s", path.c_str(), strerror(errno));
    int rc = kOk;
    std::vector<char> staging;
    for (auto& t : img->tensors) {
        const size_t nbytes = t.end - t.begin;
        if (t.dtype == "F16") {
            if (nbytes != t.numel * 2) {
                rc = fail(kErrIo, "%s: tensor %s size mismatch", path.c_str(), t.name.c_str());
                break;
            }
            if (parallel_pread(fd, img->pinned + t.host_off, data_start + t.begin, nbytes)) {
                rc = fail(kErrIo, "%s: short read of tensor %s", path.c_str(), t.name.c_str());
                break;
            }
        } else {
            staging.resize(nbytes);
            if (parallel_pread(fd, staging.data(), data_start + t.begin, nbytes)) {
                rc = fail(kErrIo, "%s: short read of tensor %s", path.c_str(), t.name.c_str());
                break;
            }
            convert_to_f16(staging.data(), t.dtype, t.numel, reinterpret_cast<__half*>(img->pinned + t.host_off));
        }
        g_file_bytes.fetch_add(nbytes);
    }
    // the pinned copy is the resident one: let the kernel drop the file's pages instead of keeping the model in RAM
    // twice (reference utils.py:72-73,126-127 reads the whole file into a bytes object first)
    posix_fadvise(fd, 0, 0, POSIX_FADV_DONTNEED);
    close(fd);
    if (rc != kOk) return rc;

    {
        std::lock_guard<std::mutex> lk(g_mu);
        auto& slot = g_cache[path];
        if (!slot.img) slot.img = img;
        slot.last_use = ++g_clock;
        *out = slot.img;
    }
    img.reset();
    if (const size_t lim = g_limit.load()) trim_host_cache(lim);  // `*out` is in use, so it survives
    return kOk;
}

void set_host_cache_limit(size_t max_bytes) {
    g_limit.store(max_bytes);
    if (max_bytes) trim_host_cache(max_bytes);
}

size_t trim_host_cache(size_t keep_bytes) {
    std::lock_guard<std::mutex> lk(g_mu);
    size_t held = 0;
    for (auto& kv : g_cache) held += kv.second.img->pinned_bytes;
    while (held > keep_bytes) {
        auto victim = g_cache.end();
        for (auto it = g_cache.begin(); it != g_cache.end(); ++it)
            if (it->second.img.use_count() == 1 && (victim == g_cache.end() || it->second.last_use < victim->second.last_use))
                victim = it;
        if (victim == g_cache.end()) break;
        held -= victim->second.img->pinned_bytes;
        g_cache.erase(victim);
    }
    return held;
}

uint64_t host_cache_file_bytes() { return g_file_bytes.load(); }
uint64_t host_cache_pinned_bytes() { return g_pinned_bytes.load(); }
uint64_t host_cache_mapped_bytes() { return g_mapped_bytes.load(); }
void set_stage_dir(const char* dir) {
    std::lock_guard<std::mutex> lk(g_stage_mu);
    g_stage_dir = dir ? dir : "";
}
void set_host_image_mode(int mode) { g_image_mode.store(mode < 0 || mode > 2 ? 0 : mode); }

}  // namespace fls
