// dh_noalloc.cu — dh_bgzf_inflate_noalloc: the BGZF inflate kernel for callers that may not allocate (code under CUDA graph
// capture, tight serving loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
// It is the lane-group kernel of bgzf_inflate.cu (bgzf_inflate_kernel: inflate + CRC-32 + ISIZE per block) with
// everything the other entry points keep in cudaMalloc'ed buffers — block descriptors, per-block verdicts, the work ticket, the
// first-failure pair — in ONE file-scope __device__ array of fixed size.  Nothing here calls cudaMalloc / cudaFree / cudaMallocHost
// or a routine that does; the host-side staging of the descriptors is a static array too.
//
// The array's name is a host-side stand-in in host code, not a device address: the device address comes from
// cudaGetSymbolAddress, asked once per device with that device current (g_scratch_ptr below), and is what the kernels get.
// Every call on a device shares the array and the array keeps the last call's contents: the verdict words are re-initialised by
// every call, the ticket is zeroed by the launcher, and calls are serialised by a process-wide mutex that is held until the
// stream has drained — so two calls from one process never overlap on the GPU.  (Two PROCESSES each have their own copy.)
#include <mutex>

#include "dh_internal.h"

namespace {

constexpr uint32_t MAX_BLOCKS = DH_NOALLOC_MAX_BLOCKS;
constexpr size_t OFF_COUNTER = 0;                                  // unsigned int: work ticket of the kernel
constexpr size_t OFF_VERDICT = 64;                                 // int[2]: {first failing block or -1, its code}
constexpr size_t OFF_STATUS = 256;                                 // int[MAX_BLOCKS]
constexpr size_t OFF_DESC = OFF_STATUS + (size_t)MAX_BLOCKS * sizeof(int);   // dh::BlockDesc[MAX_BLOCKS]
constexpr size_t SCRATCH_BYTES = OFF_DESC + (size_t)MAX_BLOCKS * sizeof(dh::BlockDesc);
static_assert(OFF_DESC % 256 == 0, "descriptor table must stay 256-byte aligned");
static_assert(SCRATCH_BYTES == DH_NOALLOC_SCRATCH_BYTES, "keep the header's figure in step");

// 590,080 bytes on every GPU that loads the library, until the process exits
static __device__ __align__(256) unsigned char g_noalloc_scratch[SCRATCH_BYTES];

constexpr int MAX_DEV = 64;
std::mutex g_mu;
unsigned char *g_scratch_ptr[MAX_DEV] = {};   // device address of g_noalloc_scratch on each device (cudaGetSymbolAddress)
int g_sm_count[MAX_DEV] = {};
dh::BlockDesc g_host_desc[MAX_BLOCKS];        // host staging of the descriptors (static: no allocation per call)

}  // namespace

extern "C" int dh_bgzf_inflate_noalloc(int device, const uint8_t *d_comp, uint64_t comp_cap, const dh_bgzf_block *blocks, uint32_t n_blocks,
                                       uint8_t *d_out, uint64_t out_cap, uint64_t *out_size, int64_t *bad_block, int32_t *bad_code,
                                       void *stream) {
    using namespace dh;
    if (out_size) *out_size = 0;
    if (bad_block) *bad_block = -1;
    if (bad_code) *bad_code = 0;
    if (device < 0 || device >= MAX_DEV) { set_error("dh_bgzf_inflate_noalloc: device %d out of range", device); return DH_EINVAL; }
    if (n_blocks > MAX_BLOCKS) {
        set_error("dh_bgzf_inflate_noalloc: %u blocks exceed the fixed scratch of DH_NOALLOC_MAX_BLOCKS = %u blocks per call "
                  "(split the call, or use dh_bgzf_inflate_host / dh_bgzf_dev_*)", n_blocks, MAX_BLOCKS);
        return DH_EINVAL;
    }
    if (n_blocks == 0) return DH_OK;
    if (!d_comp || !blocks || !d_out) { set_error("dh_bgzf_inflate_noalloc: NULL buffer"); return DH_EINVAL; }
    if (((uintptr_t)d_comp & 15u) != 0) { set_error("dh_bgzf_inflate_noalloc: d_comp must be 16-byte aligned"); return DH_EINVAL; }

    std::lock_guard<std::mutex> lk(g_mu);
    // descriptors: block i's DEFLATE payload sits at d_comp + coffset + 18, its bytes go to d_out + (sum of ISIZE before it)
    uint64_t uoff = 0;
    for (uint32_t i = 0; i < n_blocks; ++i) {
        const dh_bgzf_block &b = blocks[i];
        if (b.csize < 26 || b.isize > 65536u || b.coffset + b.csize + DH_COMP_PAD > comp_cap || b.coffset + 18 > 0xffffffffull) {
            set_error("dh_bgzf_inflate_noalloc: block %u (offset %llu, %u bytes) does not fit d_comp[0..%llu) with %u readable bytes behind it",
                      i, (unsigned long long)b.coffset, b.csize, (unsigned long long)comp_cap, DH_COMP_PAD);
            return DH_EINVAL;
        }
        BlockDesc &o = g_host_desc[i];
        o.in_off = (uint32_t)(b.coffset + 18);
        o.in_len = b.csize - 26;
        o.isize = b.isize;
        o.crc = b.crc;
        o.out_off = uoff;
        o.cmd_off = 0;
        uoff += b.isize;
    }
    if (uoff > out_cap) { set_error("dh_bgzf_inflate_noalloc: output needs %llu bytes, out_cap is %llu", (unsigned long long)uoff, (unsigned long long)out_cap); return DH_EINVAL; }

    int prev = 0;
    DH_CUDA(cudaGetDevice(&prev));
    if (prev != device) DH_CUDA(cudaSetDevice(device));
    struct Restore {
        int prev, dev;
        ~Restore() { if (prev != dev) cudaSetDevice(prev); }
    } restore{prev, device};
    if (!g_scratch_ptr[device]) {   // once per device, with that device current
        void *p = nullptr;
        DH_CUDA(cudaGetSymbolAddress(&p, g_noalloc_scratch));
        int sms = 0;
        DH_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device));
        g_sm_count[device] = sms;
        g_scratch_ptr[device] = (unsigned char *)p;
    }
    unsigned char *sc = g_scratch_ptr[device];
    cudaStream_t s = (cudaStream_t)stream;
    unsigned int *counter = (unsigned int *)(sc + OFF_COUNTER);
    int *verdict = (int *)(sc + OFF_VERDICT);
    int *status = (int *)(sc + OFF_STATUS);
    BlockDesc *desc = (BlockDesc *)(sc + OFF_DESC);

    // the array holds the last call's contents: a block the kernel does not reach must not read as "ok"
    DH_CUDA(cudaMemsetAsync(sc, 0xff, OFF_STATUS + (size_t)n_blocks * sizeof(int), s));
    DH_CUDA(cudaMemcpyAsync(desc, g_host_desc, (size_t)n_blocks * sizeof(BlockDesc), cudaMemcpyHostToDevice, s));
    DH_CUDA(launch_bgzf_inflate(d_comp, desc, n_blocks, d_out, status, counter, g_sm_count[device], s, false));   // zeroes the ticket itself
    DH_CUDA(launch_status_first_bad(status, n_blocks, verdict, s));
    int v[2] = {0, 0};
    DH_CUDA(cudaMemcpyAsync(v, verdict, sizeof(v), cudaMemcpyDeviceToHost, s));
    DH_CUDA(cudaStreamSynchronize(s));
    if (out_size) *out_size = uoff;
    if (v[0] >= 0) {
        if (bad_block) *bad_block = v[0];
        if (bad_code) *bad_code = v[1];
        set_error("BGZF block %d failed inflate/CRC check (code %d)", v[0], v[1]);
        return DH_EDATA;
    }
    return DH_OK;
}
