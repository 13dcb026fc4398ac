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

// One layer file resident in page-locked host memory as fp16 tensors (256-byte aligned each).
struct HostLayerImage {
    std::string path;
    std::vector<TensorEntry> tensors;
    char* pinned = nullptr;
    size_t pinned_bytes = 0;
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
// Returns the cached image for `path`, reading the file on first use.
int acquire_host_image(const std::string& path, std::shared_ptr<HostLayerImage>* out);
// Evicts least-recently-used images nobody holds until at most keep_bytes remain; returns bytes held.
size_t trim_host_cache(size_t keep_bytes);
// Upper bound for the pinned bytes the cache keeps (0 = unlimited): beyond it, least-recently-used images that no
// context is copying from are dropped whenever a new one is loaded (they are re-read from their files on next use).
void set_host_cache_limit(size_t max_bytes);
uint64_t host_cache_file_bytes();
uint64_t host_cache_pinned_bytes();

}  // namespace fls
