// Warp-distributed sorted list of 128 keys (4 per lane) with bitonic sort / merge networks: the top-k container of the
// kNN kernels (voxel.cu: integer voxel grid; knn_points.cu: raw float points).
#pragma once
#include <cstdint>

// ------------------------------------------------------------------ warp bitonic top-k (128 slots, 4 per lane)
// Element e = lane * 4 + j.  Keys are (d^2 << 32) | original index, so ascending key order is the contract
// "ascending (d^2, index)".
typedef unsigned long long u64;

template <typename KT>
__device__ __forceinline__ void cswap(KT& a, KT& b, bool asc) {
  KT lo = a < b ? a : b, hi = a < b ? b : a;
  a = asc ? lo : hi;
  b = asc ? hi : lo;
}

// one compare-exchange layer of the network: partner distance `s` elements, block size `k` elements
template <typename KT>
__device__ __forceinline__ void bitonic_layer(KT (&v)[4], int lane, int k, int s) {
  if (s >= 4) {
    int ls = s >> 2;
    bool lower = (lane & ls) == 0;
    bool asc = k >= 128 ? true : ((lane & (k >> 2)) == 0);
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      KT other = __shfl_xor_sync(0xffffffffu, v[j], ls);
      KT mn = v[j] < other ? v[j] : other, mx = v[j] < other ? other : v[j];
      v[j] = (lower == asc) ? mn : mx;
    }
  } else if (s == 2) {
    bool asc = k >= 128 ? true : (k >= 8 ? ((lane & (k >> 2)) == 0) : true /* k == 4: block = lane */);
    if (k == 4) asc = (lane & 1) == 0;
    cswap(v[0], v[2], asc);
    cswap(v[1], v[3], asc);
  } else {
    bool asc01, asc23;
    if (k == 2) {
      asc01 = true;
      asc23 = false;
    } else if (k == 4) {
      asc01 = asc23 = (lane & 1) == 0;
    } else {
      asc01 = asc23 = k >= 128 ? true : ((lane & (k >> 2)) == 0);
    }
    cswap(v[0], v[1], asc01);
    cswap(v[2], v[3], asc23);
  }
}

template <typename KT>
__device__ __forceinline__ void bitonic_sort128(KT (&v)[4], int lane) {
#pragma unroll
  for (int k = 2; k <= 128; k <<= 1) {
#pragma unroll
    for (int s = k >> 1; s > 0; s >>= 1) bitonic_layer(v, lane, k, s);
  }
}

template <typename KT>
__device__ __forceinline__ void bitonic_merge128(KT (&v)[4], int lane) {
#pragma unroll
  for (int s = 64; s > 0; s >>= 1) bitonic_layer(v, lane, 128, s);
}

template <typename KT>
__device__ __forceinline__ KT topk_element(const KT (&v)[4], int e) {
  int j = e & 3;
  KT x = j == 0 ? v[0] : (j == 1 ? v[1] : (j == 2 ? v[2] : v[3]));
  return __shfl_sync(0xffffffffu, x, e >> 2);
}

// merge `nbuf` buffered keys (shared memory) into the sorted list
template <typename KT>
__device__ __forceinline__ void topk_flush(KT (&list)[4], const KT* buf, int nbuf, int lane) {
#ifdef GP_TOPK_SMALL
  if (nbuf <= 32) {
    // small batch (end-of-shell leftovers): sort one key per lane (15 compare-exchange steps on a single element
    // instead of the 28 four-element layers of the 128-key network), then merge.  With the batch sorted ascending in
    // slots 0..31 of a 128-slot array padded with KMAX, the partner of list element e in the bitonic merge is slot
    // 127 - e, which is a real key only for e >= 96.
    KT x = lane < nbuf ? buf[lane] : (KT)~(KT)0;
#pragma unroll
    for (int k = 2; k <= 32; k <<= 1) {
#pragma unroll
      for (int s = k >> 1; s > 0; s >>= 1) {
        const KT other = __shfl_xor_sync(0xffffffffu, x, s);
        const bool lower = (lane & s) == 0;
        const bool asc = k == 32 ? true : ((lane & k) == 0);
        const KT mn = x < other ? x : other, mx = x < other ? other : x;
        x = (lower == asc) ? mn : mx;
      }
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int e = lane * 4 + j;
      const KT rev = __shfl_sync(0xffffffffu, x, (127 - e) & 31);
      if (e >= 96) list[j] = list[j] < rev ? list[j] : rev;
    }
    bitonic_merge128(list, lane);
    return;
  }
#endif
  KT b[4];
#pragma unroll
  for (int j = 0; j < 4; ++j) b[j] = (lane * 4 + j < nbuf) ? buf[lane * 4 + j] : (KT)~(KT)0;
  bitonic_sort128(b, lane);
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    KT rev = __shfl_sync(0xffffffffu, b[3 - j], 31 - lane);  // element 127 - e
    list[j] = list[j] < rev ? list[j] : rev;
  }
  bitonic_merge128(list, lane);
}
