// source_select.h -- how the gunrock.h entry points choose a source vertex.
// Follows the reference exactly so its golden outputs hold:
//   randomize       graphio::RandomNode (gunrock/graphio/utils.cuh:38-45) on top of
//                   util::RandomBits (gunrock/util/random_bits.h:45-70): one byte
//                   (rand() >> 7) per key byte from UNSEEDED libc rand(), |key| % n
//   largest_degree  Csr::GetNodeWithHighestDegree (gunrock/csr.cuh:442-455): first
//                   vertex of strictly largest out-degree
//   manually        GunrockConfig.src_node
#pragma once
#include <cstdlib>
#include <cstring>
#include "../../../include/gunrock/gunrock.h"

static inline int grb_random_node(long long n)
{
    unsigned char b[4];
    for (int j = 0; j < 4; ++j) b[j] = (unsigned char)(0xff & (rand() >> 7));
    int key;
    memcpy(&key, b, 4);
    if (key < 0) key *= -1;
    return n > 0 ? (int)(key % n) : 0;
}

static inline int grb_pick_source(enum SrcMode mode, int src_node, long long n, const int *row_offsets,
                                  int fallback = 0)
{
    switch (mode) {
        case randomize: return grb_random_node(n);
        case largest_degree: {
            int best = 0, src = 0;
            for (long long v = 0; v < n; ++v) {
                int d = row_offsets[v + 1] - row_offsets[v];
                if (d > best) { best = d; src = (int)v; }
            }
            return src;
        }
        case manually: return src_node;
        default: return fallback;
    }
}
