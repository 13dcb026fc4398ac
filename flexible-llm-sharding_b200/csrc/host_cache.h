#pragma once
#include <cstddef>
#include <cstdint>
#include <memory>
#include <string>
#include <vector>

namespace fls {

struct TensorEntry {
    std::string name;
    std::string dtype;
    std::vector<long long> shape;
    size_t begin = 0, end = 0;  // data_offsets relative to the data section of the file
    size_t numel = 0;
    size_t host_off = 0;  // offset of the fp16 image of this tensor inside HostLayerImage::pinned
};

// One layer file resident in page-locked host memory as fp16 tensors.  Two forms:
//   mapped  the file's own pages (tmpfs: the driver locks file pages of no other file system), mmap'ed and page-locked in place with
//           cudaHostRegister: no second copy in RAM, and every process that maps the file pins the SAME physical
//           pages - one host copy of a layer serves every rank of the box (reference utils.py:35-41,67-75 does this
//           for the threads of one process).  Needs all tensors stored as F16.
//   copied  a cudaHostAlloc'ed buffer the tensors are read (and, for BF16/F32 files, converted) into, each
//           256-byte aligned; the file's pages are then dropped from the page cache.
struct HostLayerImage {
    std::string path;
    std::vector<TensorEntry> tensors;
    char* pinned = nullptr;   // base the tensors' host_off are relative to
    size_t pinned_bytes = 0;
    void* map_base = nullptr;  // mapped form: the mmap'ed (and registered) range
    size_t map_len = 0;
    ~HostLayerImage();
    const TensorEntry* find(const std::string& suffix) const {
        for (auto& t : tensors)
            if (t.name.size() >= suffix.size() && t.name.compare(t.name.size() - suffix.size(), suffix.size(), suffix) == 0)
                return &t;
        return nullptr;
    }
};

int parse_safetensors_header(const std::string& path, std::vector<TensorEntry>* out, size_t* data_start,
                             size_t* file_size);
// `<layer>.flspack` (prepare_weights.py --packed) or `<layer>.safetensors`, by extension
int parse_layer_header(const std::string& path, std::vector<TensorEntry>* out, size_t* data_start, size_t* file_size);
// Returns the cached image for `path`, reading the file on first use.
int acquire_host_image(const std::string& path, std::shared_ptr<HostLayerImage>* out);
// Evicts least-recently-used images nobody holds until at most keep_bytes remain; returns bytes held.
size_t trim_host_cache(size_t keep_bytes);
// Upper bound for the pinned bytes the cache keeps (0 = unlimited): beyond it, least-recently-used images that no
// context is copying from are dropped whenever a new one is loaded (they are re-read from their files on next use).
void set_host_cache_limit(size_t max_bytes);
uint64_t host_cache_file_bytes();
uint64_t host_cache_pinned_bytes();
uint64_t host_cache_mapped_bytes();  // part of pinned_bytes that is file pages locked in place (no copy)
// 0 = auto (map when the file allows it and the driver accepts, else copy), 1 = always copy, 2 = map or fail
void set_host_image_mode(int mode);
// Directory on tmpfs in which disk-resident F16 layer files are staged once per box and then locked in place by every
// process (nullptr / "" = off).
void set_stage_dir(const char* dir);

}  // namespace fls
