// Which (mmap flags, cudaHostRegister flags) combinations does this driver accept for a file's own pages?
// Build: nvcc -o hostreg_probe tools/hostreg_probe.cu ; run: ./hostreg_probe <file> [<file> ...]
#include <cuda_runtime.h>
#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>
#include <chrono>
#include <cstdio>
#include <cstring>
#include <cerrno>

int main(int argc, char** argv) {
    cudaFree(0);
    int ro = 0;
    cudaDeviceGetAttribute(&ro, cudaDevAttrHostRegisterReadOnlySupported, 0);
    printf("cudaDevAttrHostRegisterReadOnlySupported = %d\n", ro);
    for (int a = 1; a < argc; ++a) {
        struct {
            const char* name;
            int oflag, prot, mflag;
            unsigned rflag;
        } combos[] = {
            {"O_RDONLY PROT_READ MAP_SHARED  + ReadOnly", O_RDONLY, PROT_READ, MAP_SHARED, cudaHostRegisterPortable | cudaHostRegisterReadOnly},
            {"O_RDONLY PROT_READ MAP_PRIVATE + ReadOnly", O_RDONLY, PROT_READ, MAP_PRIVATE, cudaHostRegisterPortable | cudaHostRegisterReadOnly},
            {"O_RDONLY PROT_READ MAP_SHARED  (default)", O_RDONLY, PROT_READ, MAP_SHARED, cudaHostRegisterPortable},
            {"O_RDWR   PROT_RW   MAP_SHARED  (default)", O_RDWR, PROT_READ | PROT_WRITE, MAP_SHARED, cudaHostRegisterPortable},
            {"O_RDWR   PROT_RW   MAP_SHARED  + ReadOnly", O_RDWR, PROT_READ | PROT_WRITE, MAP_SHARED, cudaHostRegisterPortable | cudaHostRegisterReadOnly},
            {"O_RDONLY PROT_RW   MAP_PRIVATE (default)", O_RDONLY, PROT_READ | PROT_WRITE, MAP_PRIVATE, cudaHostRegisterPortable},
        };
        for (auto& c : combos) {
            int fd = open(argv[a], c.oflag);
            if (fd < 0) { printf("%s: %s: open failed: %s\n", argv[a], c.name, strerror(errno)); continue; }
            struct stat st; fstat(fd, &st);
            size_t len = (st.st_size + 4095) / 4096 * 4096;
            void* m = mmap(nullptr, len, c.prot, c.mflag, fd, 0);
            close(fd);
            if (m == MAP_FAILED) { printf("%s: %s: mmap failed: %s\n", argv[a], c.name, strerror(errno)); continue; }
            auto t0 = std::chrono::steady_clock::now();
            cudaError_t e = cudaHostRegister(m, len, c.rflag);
            double ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
            printf("%s (%zu MB): %-45s -> %s", argv[a], len >> 20, c.name, cudaGetErrorString(e));
            if (e == cudaSuccess) {
                void* d; cudaMalloc(&d, len);
                cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
                cudaEventRecord(e0); cudaError_t ce = cudaMemcpyAsync(d, m, len, cudaMemcpyHostToDevice, 0); cudaEventRecord(e1);
                cudaEventSynchronize(e1); float cms = 0; cudaEventElapsedTime(&cms, e0, e1);
                printf("  register %.1f ms (%.2f GB/s), H2D %s %.2f GB/s", ms, len / ms / 1e6, cudaGetErrorString(ce), len / cms / 1e6);
                cudaFree(d); cudaHostUnregister(m);
            } else cudaGetLastError();
            printf("\n");
            munmap(m, len);
        }
    }
    return 0;
}
