// Probe: does cuFile (GPUDirect Storage, or its compatibility mode) work on this box?
#include <cuda_runtime.h>
#include <cufile.h>
#include <fcntl.h>
#include <unistd.h>
#include <cstdio>
#include <cstdlib>
#include <cstring>
int main(int argc, char** argv) {
    const char* path = argc > 1 ? argv[1] : "/tmp/cufile_probe.bin";
    const size_t n = 8 << 20;
    {
        int fd = open(path, O_CREAT | O_WRONLY | O_TRUNC, 0600);
        if (fd < 0) { perror("create"); return 1; }
        char* b = (char*)malloc(n);
        for (size_t i = 0; i < n; i++) b[i] = (char)(i * 2654435761u >> 24);
        if (write(fd, b, n) != (ssize_t)n) { perror("write"); return 1; }
        fsync(fd); close(fd); free(b);
    }
    CUfileError_t st = cuFileDriverOpen();
    printf("cuFileDriverOpen: err=%d cu_err=%d\n", (int)st.err, (int)st.cu_err);
    if (st.err != CU_FILE_SUCCESS) return 2;
    int fd = open(path, O_RDONLY | O_DIRECT);
    printf("open O_DIRECT fd=%d\n", fd);
    if (fd < 0) { fd = open(path, O_RDONLY); printf("open plain fd=%d\n", fd); }
    CUfileDescr_t d; memset(&d, 0, sizeof d);
    d.handle.fd = fd; d.type = CU_FILE_HANDLE_TYPE_OPAQUE_FD;
    CUfileHandle_t h;
    st = cuFileHandleRegister(&h, &d);
    printf("cuFileHandleRegister: err=%d\n", (int)st.err);
    if (st.err != CU_FILE_SUCCESS) return 3;
    void* dptr; cudaMalloc(&dptr, n);
    st = cuFileBufRegister(dptr, n, 0);
    printf("cuFileBufRegister: err=%d\n", (int)st.err);
    ssize_t r = cuFileRead(h, dptr, n, 0, 0);
    printf("cuFileRead: %zd\n", r);
    unsigned char back[16]; cudaMemcpy(back, dptr, 16, cudaMemcpyDeviceToHost);
    printf("first bytes: %02x %02x %02x %02x\n", back[0], back[1], back[2], back[3]);
    cuFileBufDeregister(dptr); cuFileHandleDeregister(h); cuFileDriverClose();
    return r == (ssize_t)n ? 0 : 4;
}
