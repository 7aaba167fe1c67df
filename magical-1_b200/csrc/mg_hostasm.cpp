/*
 * mg_hostasm.cpp -- host half of the end-to-end step (mg_step_host with mg_set_host_assembly on).
 *
 * The observation the reference returns from env.step() under the LoRes4* wrappers is the stack of the last four
 * 96x96 RGB frames, interleaved per pixel (FlattenFrameStack, magical/benchmarks/__init__.py:122-147: u8[96,96,12]).
 * Three of the four frames of every stack are already on the host after the previous step, so for part of the batch
 * mg_step_host sends only the NEW frame over PCIe (27 648 B per env instead of 110 592 B) into a pinned mirror of the
 * device's frame ring, and the threads of this file build the caller's stacked observation from the mirror while
 * the rest of the batch arrives as finished stacks.  Same bytes as the device-side stack (tests/test_gpu_hostasm.py).
 *
 * No CUDA in this file: the workers only touch host memory.  One pool per handle, started on first use.
 */
#include <atomic>
#include <condition_variable>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <thread>
#include <vector>
#if defined(__x86_64__)
#include <immintrin.h>
#endif
#include "mg_kernels.h"      /* MG_MAX_RING, MG_FRAME_BYTES, MG_MAX_GROUPS */
#include "mg_hostasm.h"

namespace {

constexpr int FRAME_BYTES = MG_FRAME_BYTES;
constexpr int N_PIX = MG_LORES * MG_LORES;

/* generic: any number of groups, byte by byte */
void assemble_scalar(uint8_t *dst, const uint8_t *const *src, int G)
{
    for (int p = 0; p < N_PIX; p++)
        for (int g = 0; g < G; g++) {
            dst[(size_t)p * 3 * G + 3 * g] = src[g][3 * p];
            dst[(size_t)p * 3 * G + 3 * g + 1] = src[g][3 * p + 1];
            dst[(size_t)p * 3 * G + 3 * g + 2] = src[g][3 * p + 2];
        }
}

#if defined(__x86_64__)
/* four groups, four pixels per round: 12 bytes of each of the four frames -> 48 interleaved bytes.  Output byte J of the
 * 48 = byte 3 * (J / 12) + J % 3 of frame (J % 12) / 3: three byte shuffles per frame, OR-ed. */
struct Masks { alignas(16) uint8_t m[3][4][16]; };
const Masks &masks()
{
    static const Masks M = [] {
        Masks t;
        for (int k = 0; k < 3; k++)
            for (int g = 0; g < 4; g++)
                for (int j = 0; j < 16; j++) {
                    const int J = 16 * k + j, px = J / 12, within = J % 12;
                    t.m[k][g][j] = within / 3 == g ? (uint8_t)(3 * px + within % 3) : (uint8_t)0x80;
                }
        return t;
    }();
    return M;
}

__attribute__((target("ssse3")))
void assemble4_ssse3(uint8_t *dst, const uint8_t *const *src, bool stream)
{
    const Masks &M = masks();
    __m128i mk[3][4];
    for (int k = 0; k < 3; k++)
        for (int g = 0; g < 4; g++) mk[k][g] = _mm_load_si128((const __m128i *)M.m[k][g]);
    const uint8_t *a = src[0], *b = src[1], *c = src[2], *d = src[3];
    /* the 16-byte loads run 4 bytes past the 12 they use: the last round of a frame is done from a copy */
    const int rounds = N_PIX / 4;
    for (int r = 0; r < rounds; r++) {
        __m128i va, vb, vc, vd;
        if (r < rounds - 1) {
            va = _mm_loadu_si128((const __m128i *)(a + 12 * r)); vb = _mm_loadu_si128((const __m128i *)(b + 12 * r));
            vc = _mm_loadu_si128((const __m128i *)(c + 12 * r)); vd = _mm_loadu_si128((const __m128i *)(d + 12 * r));
        } else {
            alignas(16) uint8_t t[4][16] = {};
            memcpy(t[0], a + 12 * r, 12); memcpy(t[1], b + 12 * r, 12); memcpy(t[2], c + 12 * r, 12); memcpy(t[3], d + 12 * r, 12);
            va = _mm_load_si128((const __m128i *)t[0]); vb = _mm_load_si128((const __m128i *)t[1]);
            vc = _mm_load_si128((const __m128i *)t[2]); vd = _mm_load_si128((const __m128i *)t[3]);
        }
        for (int k = 0; k < 3; k++) {
            const __m128i o = _mm_or_si128(_mm_or_si128(_mm_shuffle_epi8(va, mk[k][0]), _mm_shuffle_epi8(vb, mk[k][1])),
                                           _mm_or_si128(_mm_shuffle_epi8(vc, mk[k][2]), _mm_shuffle_epi8(vd, mk[k][3])));
            __m128i *o_ptr = (__m128i *)(dst + 48 * (size_t)r + 16 * k);
            if (stream) _mm_stream_si128(o_ptr, o); else _mm_storeu_si128(o_ptr, o);
        }
    }
    if (stream) _mm_sfence();
}

/* the same with AVX-512 VBMI, sixteen pixels per round: 48 bytes of each frame -> 192 interleaved bytes.  Each 64-byte
 * output is two two-source byte permutes (frames 0|1 and 2|3) and a blend. */
struct Masks512 { alignas(64) uint8_t idx_ab[3][64], idx_cd[3][64]; uint64_t take_cd[3]; };
const Masks512 &masks512()
{
    static const Masks512 M = [] {
        Masks512 t;
        for (int k = 0; k < 3; k++) {
            t.take_cd[k] = 0;
            for (int j = 0; j < 64; j++) {
                const int J = 64 * k + j, px = J / 12, within = J % 12, g = within / 3, byte = 3 * px + within % 3;
                t.idx_ab[k][j] = (uint8_t)(byte + 64 * (g == 1));          /* table = frame 0 (0..63) | frame 1 (64..127) */
                t.idx_cd[k][j] = (uint8_t)(byte + 64 * (g == 3));
                if (g >= 2) t.take_cd[k] |= 1ull << j;
            }
        }
        return t;
    }();
    return M;
}

__attribute__((target("avx512f,avx512bw,avx512vbmi")))
void assemble4_avx512(uint8_t *dst, const uint8_t *const *src, bool stream)
{
    const Masks512 &M = masks512();
    __m512i iab[3], icd[3];
    __mmask64 kcd[3];
    for (int k = 0; k < 3; k++) {
        iab[k] = _mm512_load_si512(M.idx_ab[k]); icd[k] = _mm512_load_si512(M.idx_cd[k]);
        kcd[k] = (__mmask64)M.take_cd[k];
    }
    const __mmask64 k48 = 0xffffffffffffull;        /* 48 of the 64 bytes: masked-off bytes are not touched */
    const uint8_t *a = src[0], *b = src[1], *c = src[2], *d = src[3];
    for (int r = 0; r < N_PIX / 16; r++) {
        const __m512i va = _mm512_maskz_loadu_epi8(k48, a + 48 * r), vb = _mm512_maskz_loadu_epi8(k48, b + 48 * r);
        const __m512i vc = _mm512_maskz_loadu_epi8(k48, c + 48 * r), vd = _mm512_maskz_loadu_epi8(k48, d + 48 * r);
        for (int k = 0; k < 3; k++) {
            const __m512i ab = _mm512_permutex2var_epi8(va, iab[k], vb), cd = _mm512_permutex2var_epi8(vc, icd[k], vd);
            const __m512i o = _mm512_mask_blend_epi8(kcd[k], ab, cd);
            void *o_ptr = dst + 192 * (size_t)r + 64 * k;
            if (stream) _mm512_stream_si512((__m512i *)o_ptr, o); else _mm512_storeu_si512(o_ptr, o);
        }
    }
    if (stream) _mm_sfence();
}
#endif

}  // namespace

struct MgHostAsm {
    std::vector<std::thread> threads;
    std::mutex mu;
    std::condition_variable cv;
    uint64_t generation = 0;
    bool quit = false;
    /* the step in flight */
    const uint8_t *ring = nullptr;
    uint8_t *obs = nullptr;
    const uint8_t *since_reset = nullptr;
    int n_envs = 0, newest = 0, G = 0;
    int age[MG_MAX_GROUPS] = { 0 };
    bool ssse3 = false, avx512 = false;
    std::atomic<int> next{0}, limit{0}, done{0};

    void one_env(int e)
    {
        const uint8_t *src[MG_MAX_GROUPS];
        for (int g = 0; g < G; g++) {
            /* a frame older than the env's episode does not exist: the stack repeats the first frame
             * (FlattenFrameStack.reset, benchmarks/__init__.py:139-147) */
            const int a = age[g] < since_reset[e] ? age[g] : since_reset[e];
            const int slot = (newest - a) & (MG_MAX_RING - 1);
            src[g] = ring + ((size_t)e * MG_MAX_RING + slot) * FRAME_BYTES;
        }
        uint8_t *dst = obs + (size_t)e * FRAME_BYTES * G;
#if defined(__x86_64__)
        if (G == 4 && avx512) { assemble4_avx512(dst, src, ((uintptr_t)dst & 63) == 0); return; }
        if (G == 4 && ssse3) { assemble4_ssse3(dst, src, ((uintptr_t)dst & 15) == 0); return; }
#endif
        assemble_scalar(dst, src, G);
    }
    /* take envs until the step is through; envs beyond `limit` have not landed yet */
    void work()
    {
        for (;;) {
            const int i = next.fetch_add(1, std::memory_order_acq_rel);
            if (i >= n_envs) return;
            if (i >= limit.load(std::memory_order_acquire)) {
                /* its frame has not landed: sleep until the next publish (no spinning: the cores belong to whoever
                 * else runs on the host, e.g. the thread that samples the coming episodes) */
                std::unique_lock<std::mutex> lk(mu);
                cv.wait(lk, [&] { return i < limit.load(std::memory_order_acquire); });
            }
            one_env(i);
            done.fetch_add(1, std::memory_order_release);
        }
    }
    void loop()
    {
        uint64_t seen = 0;
        for (;;) {
            {
                std::unique_lock<std::mutex> lk(mu);
                cv.wait(lk, [&] { return quit || generation != seen; });
                if (quit) return;
                seen = generation;
            }
            work();
        }
    }
};

MgHostAsm *mg_hostasm_create(int n_threads)
{
    MgHostAsm *A = new MgHostAsm();
#if defined(__x86_64__)
    A->ssse3 = __builtin_cpu_supports("ssse3");
    A->avx512 = __builtin_cpu_supports("avx512f") && __builtin_cpu_supports("avx512bw") && __builtin_cpu_supports("avx512vbmi");
    if (const char *ev = getenv("MG_HOST_SIMD")) {       /* 0: bytes, 1: SSSE3, 2: AVX-512 (testing) */
        const int lvl = atoi(ev);
        A->avx512 = A->avx512 && lvl >= 2; A->ssse3 = A->ssse3 && lvl >= 1;
    }
#endif
    if (n_threads < 1) n_threads = 1;
    for (int t = 0; t < n_threads - 1; t++) A->threads.emplace_back([A] { A->loop(); });   /* the caller is the last worker */
    return A;
}

void mg_hostasm_destroy(MgHostAsm *A)
{
    if (!A) return;
    { std::lock_guard<std::mutex> lk(A->mu); A->quit = true; }
    A->cv.notify_all();
    for (auto &t : A->threads) t.join();
    delete A;
}

void mg_hostasm_begin(MgHostAsm *A, const uint8_t *ring_host, uint8_t *obs_host, int n_envs, int newest_slot,
                      const int *ages, int n_groups, const uint8_t *since_reset)
{
    A->ring = ring_host; A->obs = obs_host; A->n_envs = n_envs; A->newest = newest_slot; A->G = n_groups;
    for (int g = 0; g < n_groups; g++) A->age[g] = ages[g];
    A->since_reset = since_reset;
    A->next.store(0); A->limit.store(0); A->done.store(0);
    { std::lock_guard<std::mutex> lk(A->mu); A->generation++; }
    A->cv.notify_all();
}

void mg_hostasm_publish(MgHostAsm *A, int env_limit)
{
    { std::lock_guard<std::mutex> lk(A->mu); A->limit.store(env_limit, std::memory_order_release); }
    A->cv.notify_all();
}

void mg_hostasm_finish(MgHostAsm *A)
{
    A->work();
    while (A->done.load(std::memory_order_acquire) < A->n_envs) std::this_thread::yield();     /* the last few envs in flight */
}

int mg_hostasm_threads(const MgHostAsm *A) { return (int)A->threads.size() + 1; }
