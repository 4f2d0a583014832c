// splat_sort_pairs_u64_u32_static: the stand-alone (uint64 key, uint32 value) radix sort for callers that cannot allocate.
// No counterpart in the reference; added for code under CUDA graph capture, tight
// serving loops and harnesses that forbid cudaMalloc. Same kernels as splat_sort_pairs_u64_u32 (k_sort.cu); what differs
// is where their scratch lives: the control block (digit histograms, tickets) and the decoupled-lookback status words sit
// in ONE file-scope device array of a fixed size instead of buffers the context allocates on demand.
//
//   - every GPU that loads this library holds the array (STATIC_SCRATCH_BYTES, about 1 MiB) until the process exits;
//   - sorts of more than SPLAT_STATIC_SORT_MAX_PAIRS pairs are refused (SPLAT_ERR_ARG);
//   - every call on a device shares the array and finds the previous call's contents in it: the part a sort needs zeroed
//     is cleared on the caller's stream first; two calls whose work OVERLAPS on the same GPU (different streams) race;
//   - this entry point calls no cudaMalloc / cudaFree / allocating library routine.
#include "splat_internal.cuh"

#include <mutex>
#include <string>

#define STATIC_SORT_TILE (SORT_THREADS * SORT_ITEMS)                                   // 4096 pairs per onesweep tile (64-bit keys)
#define STATIC_SORT_TILES (SPLAT_STATIC_SORT_MAX_PAIRS / STATIC_SORT_TILE)             // 512
#define STATIC_CTL_BYTES ((sizeof(SortCtl) + 255) / 256 * 256)                         // control block, padded to 256 B
#define STATIC_STATUS_BYTES ((size_t)STATIC_SORT_TILES * 256 * 8)                      // one 64-bit lookback word per (tile, digit)
#define STATIC_SCRATCH_BYTES (STATIC_CTL_BYTES + STATIC_STATUS_BYTES)
static_assert(SPLAT_STATIC_SORT_MAX_PAIRS % STATIC_SORT_TILE == 0, "the maximum is a whole number of tiles");

// The scratch. In HOST code this name is only a stand-in for the symbol (not a device address): the address the kernels
// get comes from cudaGetSymbolAddress below, taken once per device with that device current.
static __device__ __align__(256) unsigned char g_static_sort_scratch[STATIC_SCRATCH_BYTES];

void splat_set_last_error(const char* msg);        // splat_api.cu: the thread's splat_last_error_string()

// Zeroes the control block (tickets and histograms must start at zero) and sets the pair count, in one launch.
__global__ void __launch_bounds__(256) static_sort_init_kernel(SortCtl* ctl, unsigned int n)
{
    unsigned int* w = reinterpret_cast<unsigned int*>(ctl);
    for (unsigned int i = threadIdx.x; i < sizeof(SortCtl) / 4; i += 256) w[i] = 0u;
    __syncthreads();
    if (threadIdx.x == 0) ctl->n_sort = n;
}

#define STATIC_MAX_DEVICES 64
struct StaticSortDevice { unsigned char* scratch; int sm_count; };
static StaticSortDevice g_dev[STATIC_MAX_DEVICES];
static std::mutex g_dev_mutex;

// Set-up, once per device, with that device current: the array's device address and the SM count. Allocates nothing.
static int static_sort_device(int device, StaticSortDevice& out)
{
    std::lock_guard<std::mutex> lock(g_dev_mutex);
    if (!g_dev[device].scratch) {
        void* p = nullptr;
        cudaError_t e = cudaGetSymbolAddress(&p, g_static_sort_scratch);
        if (e != cudaSuccess || !p) {
            splat_set_last_error((std::string("splat_sort_pairs_u64_u32_static: cudaGetSymbolAddress failed: ") + cudaGetErrorString(e)).c_str());
            return SPLAT_ERR_CUDA;
        }
        int sms = 0;
        e = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
        if (e != cudaSuccess || sms <= 0) {
            splat_set_last_error((std::string("splat_sort_pairs_u64_u32_static: cudaDeviceGetAttribute failed: ") + cudaGetErrorString(e)).c_str());
            return SPLAT_ERR_CUDA;
        }
        g_dev[device].sm_count = sms;
        g_dev[device].scratch = static_cast<unsigned char*>(p);
    }
    out = g_dev[device];
    return SPLAT_OK;
}

extern "C" int splat_sort_pairs_u64_u32_static(void* stream, uint64_t* keys, uint32_t* vals, uint64_t* keys_alt, uint32_t* vals_alt,
                                               uint64_t n, int begin_bit, int end_bit, int* result_in_alt)
{
    if (n > (uint64_t)SPLAT_STATIC_SORT_MAX_PAIRS) {
        splat_set_last_error("splat_sort_pairs_u64_u32_static: n exceeds SPLAT_STATIC_SORT_MAX_PAIRS (2097152 pairs): the fixed scratch "
                             "array cannot hold its lookback words; use splat_sort_pairs_u64_u32");
        return SPLAT_ERR_ARG;
    }
    if (n && (!keys || !vals || !keys_alt || !vals_alt)) { splat_set_last_error("splat_sort_pairs_u64_u32_static: NULL buffer"); return SPLAT_ERR_ARG; }
    if (begin_bit < 0 || end_bit > 64 || begin_bit > end_bit) { splat_set_last_error("splat_sort_pairs_u64_u32_static: bad bit range"); return SPLAT_ERR_ARG; }
    SortPlan plan{};
    for (int b = begin_bit; b < end_bit; b += 8) {
        const int bits = end_bit - b < 8 ? end_bit - b : 8;
        plan.shift[plan.n_pass] = b; plan.mask[plan.n_pass] = (1u << bits) - 1u; ++plan.n_pass;
    }
    for (int k = 0; k < SPLAT_MAX_SORT_PASSES; ++k) { plan.slot[k] = k; plan.ticket[k] = k; }
    if (result_in_alt) *result_in_alt = plan.n_pass & 1;
    if (n == 0 || plan.n_pass == 0) return SPLAT_OK;

    int device = -1;
    cudaError_t e = cudaGetDevice(&device);
    if (e != cudaSuccess || device < 0 || device >= STATIC_MAX_DEVICES) {
        splat_set_last_error("splat_sort_pairs_u64_u32_static: no current CUDA device (or a device index of 64 and more)");
        return SPLAT_ERR_CUDA;
    }
    StaticSortDevice d;
    { const int rc = static_sort_device(device, d); if (rc) return rc; }
    SortCtl* ctl = reinterpret_cast<SortCtl*>(d.scratch);
    unsigned long long* status = reinterpret_cast<unsigned long long*>(d.scratch + STATIC_CTL_BYTES);
    cudaStream_t s = (cudaStream_t)stream;

    // The array keeps the last call's contents. What must start at zero: the control block, and the lookback words of the
    // tiles this sort uses (a zero word carries no flag, so no pass can mistake a stale word for a published prefix).
    const size_t tiles = (size_t)((n + STATIC_SORT_TILE - 1) / STATIC_SORT_TILE);
    e = cudaMemsetAsync(status, 0, tiles * 256 * 8, s);
    if (e != cudaSuccess) { splat_set_last_error((std::string("splat_sort_pairs_u64_u32_static: clearing the scratch failed: ") + cudaGetErrorString(e)).c_str()); return SPLAT_ERR_CUDA; }
    static_sort_init_kernel<<<1, 256, 0, s>>>(ctl, (unsigned int)n);
    launch_sort_histogram(keys, 8, n, ctl, plan, d.sm_count, s);
    g_pdl_frame = true;
    launch_radix_sort(keys, vals, keys_alt, vals_alt, 8, n, ctl, status, 17u, plan, d.sm_count, false, s);
    e = cudaGetLastError();
    if (e != cudaSuccess) { splat_set_last_error((std::string("splat_sort_pairs_u64_u32_static: launch failed: ") + cudaGetErrorString(e)).c_str()); return SPLAT_ERR_CUDA; }
    return SPLAT_OK;
}
