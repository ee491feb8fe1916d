// orbm_internal.h -- types shared by the matcher's host side (orbm_api.cu) and its kernels (k_match.cu).
#pragma once
#include <stdint.h>

// Segments of one upload: src inside the device mirror of the context's pinned arena, dst the buffer a
// search kernel reads (k_scatter, ctx_flush).
#define ORBM_SCATTER_MAX 8
struct OrbmScatter {
    struct Seg {
        void* dst;
        const void* src;
        uint32_t bytes;
        uint32_t _pad;
    } seg[ORBM_SCATTER_MAX];
    int n;
};
