// Huffman codebook construction on the GPU: one CTA, everything resident in
// shared memory (<= 8192-entry dictionaries) or in an L2-resident global
// workspace (larger ones).
//
// reference: Lossless/ParallelHuffman/GetCodebook.hpp:23-145
//   stable SortByKey(freq, symbol)      -> bitonic sort on (freq<<16 | symbol)
//   GenerateCL (GenerateCL.hpp:30-733)  -> phase-by-phase restatement; the
//       reference's 32-thread blocks of the merge-path search map onto warps
//   GenerateCW (GenerateCW.hpp:25-250)  -> phase-by-phase restatement
//   ReverseArray + ReorderByIndex       -> final gather
// The code lengths produced by GenerateCL are not always the optimal Huffman
// lengths, so the step sequence is kept literally (see oracle/ for the CPU
// restatement this kernel is tested against, bit for bit).
#include "mgx_device.cuh"
#include "mgx_internal.h"

namespace {

#ifndef MGX_CB_THREADS
#define MGX_CB_THREADS 1024
#endif
constexpr int CB_THREADS = MGX_CB_THREADS;
constexpr int CB_WARPS = CB_THREADS / 32;

struct CbState {
  int iFront, iRear, lCur, iSize, curLeaves, tempLength, mergeFront, mergeRear;
  unsigned minFreq;
  int lCurOld; // lNodesCur before Operation5 advanced it (base of "copy")
  int n, fnz;
  // GenerateCW status
  int CCL, CDPI, newCDPI, updateEnd, curEntryVal, numCCL;
  int maxCL;
};

__device__ __forceinline__ int modi(int a, int b) { return ((a % b) + b) % b; }

struct CbArrays {
  uint16_t *ssym;    // [dict] symbols sorted by ascending (freq, symbol)
  uint32_t *sfreq;   // [dict] their frequencies
  uint8_t *CL;       // [n]
  uint16_t *lLeader; // [n] 0xFFFF = none
  uint16_t *iLeader; // [n]
  uint32_t *iFreq;   // [n]
  uint32_t *tFreq;   // [n+2]
  uint16_t *tIndex;  // [n+2] bit 15 = is leaf
  uint32_t *dpi;     // [2*(grid+1)]
  unsigned long long *CW; // [n]   (aliases iFreq/tFreq, used after GenerateCL)
  unsigned long long *sortbuf; // [P] (aliases iFreq.., used before)
};

} // namespace

__global__ void __launch_bounds__(CB_THREADS, 1)
    k_codebook(const uint32_t *__restrict__ freq, int dict, int P,
               unsigned long long *__restrict__ codebook,
               unsigned long long *__restrict__ decodebook,
               unsigned char *__restrict__ gws, int use_smem,
               unsigned long long *__restrict__ scalars) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  __shared__ CbState st;
  __shared__ int s_red[CB_WARPS];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;

  // ---- carve the workspace ---------------------------------------------------
  unsigned char *base = use_smem ? smem_raw : gws;
  CbArrays A;
  size_t o = 0;
  auto carve = [&](size_t bytes) {
    unsigned char *p = base + o;
    o += (bytes + 15) & ~(size_t)15;
    return p;
  };
  A.ssym = (uint16_t *)carve((size_t)dict * 2);
  A.sfreq = (uint32_t *)carve((size_t)dict * 4);
  A.CL = (uint8_t *)carve((size_t)dict);
  A.lLeader = (uint16_t *)carve((size_t)dict * 2);
  A.iLeader = (uint16_t *)carve((size_t)dict * 2);
  A.tIndex = (uint16_t *)carve((size_t)(dict + 2) * 2);
  A.dpi = (uint32_t *)carve((size_t)(2 * (dict / 32 + 2)) * 4);
  // big aliased region: max(P*8 sort keys, dict*8 CW, iFreq + tFreq)
  unsigned char *big = carve((size_t)P * 8 + 16);
  A.sortbuf = (unsigned long long *)big;
  A.CW = (unsigned long long *)big;
  A.iFreq = (uint32_t *)big;
  A.tFreq = (uint32_t *)(big + (size_t)dict * 4);
  // tFreq needs dict+2 entries: P*8+16 >= dict*8+8 always holds

  // ---- stable sort by frequency (SortByKey, GetCodebook.hpp:40-51) ----------
  // Most of the dictionary has count 0: those symbols stay in index order at the
  // front (what a stable sort gives), only the non-zero ones are sorted, by
  // (count, index).  Step 1: ordered compaction -- a thread owns `per` consecutive
  // symbols, an exclusive scan of its non-zero count gives their ranks.
  {
    const int per = (dict + CB_THREADS - 1) / CB_THREADS;
    const int lo = min(tid * per, dict), hi = min(lo + per, dict);
    int nz = 0;
    for (int i = lo; i < hi; i++)
      nz += freq[i] != 0;
    int incl = nz;
    for (int off = 1; off < 32; off <<= 1) {
      const int t = __shfl_up_sync(0xffffffffu, incl, off);
      if (lane >= off)
        incl += t;
    }
    if (lane == 31)
      s_red[warp] = incl;
    __syncthreads();
    int before = incl - nz, total = 0;
    for (int w = 0; w < CB_WARPS; w++) {
      const int c = s_red[w];
      if (w < warp)
        before += c;
      total += c;
    }
    if (tid == 0) {
      st.n = total;
      st.fnz = dict - total;
    }
    int rz = lo - before; // zero-count symbols before this thread's range
    for (int i = lo; i < hi; i++) {
      const uint32_t f = freq[i];
      if (f) {
        A.sortbuf[before++] = ((unsigned long long)f << 16) | (unsigned)i;
      } else {
        A.ssym[rz] = (uint16_t)i;
        A.sfreq[rz] = 0;
        ++rz;
      }
    }
    __syncthreads();
  }
  {
    const int n_ = st.n, fnz_ = st.fnz;
    int P2 = 1;
    while (P2 < n_)
      P2 <<= 1;
    for (int i = n_ + tid; i < P2; i += CB_THREADS)
      A.sortbuf[i] = ~0ull;
    __syncthreads();
    for (int k = 2; k <= P2; k <<= 1)
      for (int j = k >> 1; j > 0; j >>= 1) {
        for (int i = tid; i < P2; i += CB_THREADS) {
          int ixj = i ^ j;
          if (ixj > i) {
            unsigned long long a = A.sortbuf[i], b = A.sortbuf[ixj];
            bool up = (i & k) == 0;
            if ((a > b) == up) {
              A.sortbuf[i] = b;
              A.sortbuf[ixj] = a;
            }
          }
        }
        __syncthreads();
      }
    // GetFirstNonzeroIndex (GetCodebook.hpp:53-61) = number of zero counts
    for (int i = tid; i < n_; i += CB_THREADS) {
      const unsigned long long k = A.sortbuf[i];
      A.ssym[fnz_ + i] = (uint16_t)(k & 0xffff);
      A.sfreq[fnz_ + i] = (uint32_t)(k >> 16);
    }
    __syncthreads();
  }
  const int n = st.n, fnz = st.fnz;
  if (n <= 0) {
    if (tid == 0)
      scalars[MGX_SC_ERR] |= 1ull;
    return;
  }
  const uint32_t *lf = A.sfreq + fnz; // lNodesFreq == histogram (nz view)
  uint8_t *CL = A.CL;
  const int grid = (n - 1) / 32 + 1; // the reference's block count
#define FREQ_AT(i) ((i) < n ? lf[(i)] : 0u) /* see oracle note on the OOB read */

  // ---- GenerateCL ---------------------------------------------------------------
  for (int i = tid; i < n; i += CB_THREADS) { // Operation1
    A.lLeader[i] = 0xFFFF;
    A.iLeader[i] = 0xFFFF;
    CL[i] = 0;
  }
  if (tid == 0) {
    st.iFront = st.iRear = st.lCur = st.iSize = 0;
    st.curLeaves = st.tempLength = st.mergeFront = st.mergeRear = 0;
  }
  __syncthreads();
  bool cont = (st.lCur < n || st.iSize > 1); // LoopCondition1
  __syncthreads(); // nobody may enter Operation2 before everyone has read st
  while (cont) {
    if (tid == 0) { // Operation2
      unsigned mf[4] = {0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu};
      int ml[4] = {0, 0, 0, 0};
      int lCur = st.lCur, iFront = st.iFront, iRear = st.iRear, iSize = st.iSize;
      if (lCur < n) { mf[0] = lf[lCur]; ml[0] = 1; }
      if (lCur < n - 1) { mf[1] = lf[lCur + 1]; ml[1] = 1; }
      if (iSize >= 1) { mf[2] = A.iFreq[iFront]; ml[2] = 0; }
      if (iSize >= 2) { mf[3] = A.iFreq[modi(iFront + 1, n)]; ml[3] = 0; }
#define CSWAP(a, b)                                                            \
  if (mf[a] > mf[b]) {                                                         \
    unsigned tf = mf[a]; mf[a] = mf[b]; mf[b] = tf;                            \
    int tl = ml[a]; ml[a] = ml[b]; ml[b] = tl;                                 \
  }
      CSWAP(1, 3) CSWAP(0, 2) CSWAP(0, 1) CSWAP(2, 3) CSWAP(1, 2)
#undef CSWAP
      unsigned minFreq = mf[0];
      if (mf[1] < 0xFFFFFFFFu)
        minFreq += mf[1];
      A.iFreq[iRear] = minFreq;
      A.iLeader[iRear] = 0xFFFF;
      for (int k = 0; k < 2; k++)
        if (mf[k] < 0xFFFFFFFFu) {
          if (ml[k]) {
            A.lLeader[lCur] = (uint16_t)iRear;
            ++CL[lCur];
            ++lCur;
          } else {
            A.iLeader[iFront] = (uint16_t)iRear;
            iFront = modi(iFront + 1, n);
          }
        }
      st.minFreq = minFreq;
      st.lCur = lCur;
      st.iFront = iFront;
      st.iSize = modi(iRear - iFront, n);
      st.curLeaves = 0;
    }
    __syncthreads();
    { // Operation3: leaves with freq <= minFreq (sorted => a prefix count)
      const int lCur = st.lCur;
      const unsigned minFreq = st.minFreq;
      int best = 0;
      for (int i = lCur + tid; i < n; i += CB_THREADS)
        if (lf[i] <= minFreq)
          best = max(best, i - lCur + 1);
      for (int off = 16; off > 0; off >>= 1)
        best = max(best, __shfl_xor_sync(0xffffffffu, best, off));
      if (lane == 0 && best > 0)
        atomicMax(&st.curLeaves, best);
    }
    __syncthreads();
    // Operation4 (copyFreq/copyIndex) is a view: copy[j] = leaf lCurOld + j
    if (tid == 0) { // Operation5
      int iRear = st.iRear, iFront = st.iFront, iSize = st.iSize;
      int curLeaves = st.curLeaves, lCur = st.lCur;
      int mergeRear = iRear, mergeFront = iFront;
      st.lCurOld = lCur;
      if ((curLeaves + iSize) % 2 == 0) {
        iFront = iRear;
      } else if (iSize != 0 &&
                 (curLeaves == 0 ||
                  FREQ_AT(lCur + curLeaves) <= A.iFreq[modi(iRear - 1, n)])) {
        mergeRear = modi(mergeRear - 1, n);
        iFront = modi(iRear - 1, n);
      } else {
        iFront = iRear;
        --curLeaves;
      }
      lCur = lCur + curLeaves;
      iRear = modi(iRear + 1, n);
      st.tempLength = curLeaves + modi(mergeRear - mergeFront, n);
      st.curLeaves = curLeaves;
      st.lCur = lCur;
      st.iRear = iRear;
      st.iFront = iFront;
      st.mergeRear = mergeRear;
      st.mergeFront = mergeFront;
    }
    __syncthreads();
    if (st.tempLength > 0) { // BranchCondition1: merge-path partition + merge
      const int Alen = st.curLeaves;
      const int Blen = modi(st.mergeRear - st.mergeFront, n);
      const int mergeFront = st.mergeFront;
      const uint32_t *cF = lf + st.lCurOld;
      for (int b = warp; b < grid; b += CB_WARPS) { // Operation6..9, 1 warp = 1 block
        int ci = (int)(((unsigned long long)b *
                        ((unsigned long long)Alen + (unsigned long long)Blen)) /
                       (unsigned long long)grid);
        int x_top, y_top, x_bot, y_bot;
        if (Alen >= Blen) {
          x_top = min(ci, Alen);
          y_top = ci > Alen ? ci - Alen : 0;
          x_bot = y_top;
          y_bot = x_top;
        } else {
          y_bot = min(ci, Blen);
          x_bot = ci > Blen ? ci - Blen : 0;
          y_top = x_bot;
          x_top = y_bot;
        }
        bool found = false;
        int guard = 0;
        while (!found) {
          if (++guard > 64) { // never observed; keeps a bad input from hanging
            if (lane == 0) {
              A.dpi[b] = (uint32_t)x_top;
              A.dpi[b + grid + 1] = (uint32_t)y_top;
              atomicOr(&scalars[MGX_SC_ERR], 4ull);
            }
            break;
          }
          const int off = lane - 16;
          int cx = x_top - ((x_top - x_bot) >> 1) - off;
          int cy = y_top + ((y_bot - y_top) >> 1) + off;
          int gx = cx - 1, gy = mergeFront + cy;
          if (gy >= n)
            gy -= n;
          int oz;
          if (cx > Alen || cy < 0)
            oz = 0;
          else if (cy >= Blen || cx < 1)
            oz = 1;
          else
            oz = (cF[gx] <= A.iFreq[gy]) ? 1 : 0;
          unsigned m = __ballot_sync(0xffffffffu, oz);
          unsigned diff = (m ^ (m << 1)) & ~1u; // bit t: oz[t] != oz[t-1]
          if (diff) { // Operation8: the highest such thread writes last
            int tsel = 31 - __clz(diff);
            int sx = __shfl_sync(0xffffffffu, cx, tsel);
            int sy = __shfl_sync(0xffffffffu, cy, tsel);
            if (lane == 0) {
              A.dpi[b] = (uint32_t)sx;
              A.dpi[b + grid + 1] = (uint32_t)sy;
            }
            found = true;
          }
          int mx = __shfl_sync(0xffffffffu, cx, 16); // Operation9 (thread 16)
          int my = __shfl_sync(0xffffffffu, cy, 16);
          if (m >> 31) {
            x_bot = mx;
            y_bot = my;
          } else {
            x_top = mx;
            y_top = my;
          }
        }
      }
      __syncthreads();
      if (tid == 0) { // Operation10
        A.dpi[0] = 0;
        A.dpi[grid + 1] = 0;
        A.dpi[grid] = (uint32_t)Alen;
        A.dpi[grid + grid + 1] = (uint32_t)Blen;
      }
      __syncthreads();
      for (int b = tid; b < grid; b += CB_THREADS) { // Operation11
        int xs = (int)A.dpi[b], ys0 = (int)A.dpi[b + grid + 1];
        int xe = (int)A.dpi[b + 1], ye0 = (int)A.dpi[b + grid + 2];
        int ys = modi(mergeFront + ys0, n), ye = modi(mergeFront + ye0, n);
        int offset = xs + ys0, len = 0, ic = xs, ii = ys;
        const int lbase = st.lCurOld;
        while (ic < xe && modi(ye - ii, n) > 0) {
          if (cF[ic] <= A.iFreq[ii]) {
            A.tFreq[offset + len] = cF[ic];
            A.tIndex[offset + len] = (uint16_t)((lbase + ic) | 0x8000);
            ++ic;
          } else {
            A.tFreq[offset + len] = A.iFreq[ii];
            A.tIndex[offset + len] = (uint16_t)ii;
            ii = modi(ii + 1, n);
          }
          ++len;
        }
        while (ic < xe) {
          A.tFreq[offset + len] = cF[ic];
          A.tIndex[offset + len] = (uint16_t)((lbase + ic) | 0x8000);
          ++ic;
          ++len;
        }
        while (modi(ye - ii, n) > 0) {
          A.tFreq[offset + len] = A.iFreq[ii];
          A.tIndex[offset + len] = (uint16_t)ii;
          ii = modi(ii + 1, n);
          ++len;
        }
      }
      __syncthreads();
    }
    { // Operation12: meld pairs into new internal nodes
      const int half = st.tempLength / 2, iRear = st.iRear;
      // the new nodes' frequencies may overwrite iFreq entries that a later
      // pair still needs only through tFreq (already copied) -> safe
      for (int i = tid; i < half; i += CB_THREADS) {
        int ind = modi(iRear + i, n);
        A.iFreq[ind] = A.tFreq[2 * i] + A.tFreq[2 * i + 1];
        A.iLeader[ind] = 0xFFFF;
        for (int k = 0; k < 2; k++) {
          uint16_t ti = A.tIndex[2 * i + k];
          if (ti & 0x8000) {
            A.lLeader[ti & 0x7FFF] = (uint16_t)ind;
            ++CL[ti & 0x7FFF];
          } else {
            A.iLeader[ti] = (uint16_t)ind;
          }
        }
      }
    }
    __syncthreads();
    if (tid == 0) // Operation13
      st.iRear = modi(st.iRear + st.tempLength / 2, n);
    for (int i = tid; i < n; i += CB_THREADS) { // Operation14
      uint16_t ld = A.lLeader[i];
      if (ld != 0xFFFF) {
        uint16_t up = A.iLeader[ld];
        if (up != 0xFFFF) {
          A.lLeader[i] = up;
          ++CL[i];
        }
      }
    }
    __syncthreads();
    if (tid == 0) // Operation15
      st.iSize = modi(st.iRear - st.iFront, n);
    __syncthreads();
    cont = (st.lCur < n || st.iSize > 1);
    __syncthreads();
  }
#undef FREQ_AT
  if (tid == 0) {
    st.maxCL = CL[0];
    scalars[MGX_SC_MAXLEN] = CL[0];
    scalars[MGX_SC_NZ] = (unsigned long long)n;
    if (CL[0] > 56)
      scalars[MGX_SC_ERR] |= 2ull; // GetCodebook.hpp:111-119
  }
  __syncthreads();

  // ---- GenerateCW -----------------------------------------------------------------
  unsigned long long *first = decodebook, *entry = decodebook + 64,
                     *keys = decodebook + 128;
  unsigned long long *CW = A.CW;
  // workspace.reset(): decodebook = 0xff.., codebook = 0
  for (int i = tid; i < 128 + dict; i += CB_THREADS)
    decodebook[i] = ~0ull;
  for (int i = tid; i < dict; i += CB_THREADS)
    codebook[i] = 0;
  __syncthreads();
  for (int i = tid; i < n / 2; i += CB_THREADS) { // Operation1: reverse CL
    uint8_t t = CL[i];
    CL[i] = CL[n - i - 1];
    CL[n - i - 1] = t;
  }
  __syncthreads();
  if (tid == 0) { // Operation2
    st.CCL = CL[0];
    st.CDPI = 0;
    st.newCDPI = n - 1;
    entry[st.CCL] = 0;
    CW[0] = 0;
    first[st.CCL] = CW[0] ^ ((1ull << CL[0]) - 1);
    entry[st.CCL + 1] = 1;
  }
  __syncthreads();
  if (tid < st.CCL) { // Operation3
    first[tid] = ~0ull;
    entry[tid] = 0;
  }
  __syncthreads();
  while (st.CDPI < n - 1) { // LoopCondition1
    { // Operation4: first index whose successor is longer than CCL
      const int CCL = st.CCL;
      int best = 0x7fffffff;
      for (int i = tid; i < n - 1; i += CB_THREADS)
        if ((int)CL[i + 1] > CCL)
          best = min(best, i);
      for (int off = 16; off > 0; off >>= 1)
        best = min(best, __shfl_xor_sync(0xffffffffu, best, off));
      if (lane == 0 && best != 0x7fffffff)
        atomicMin(&st.newCDPI, best);
    }
    __syncthreads();
    if (tid == 0) { // Operation5
      const int newCDPI = st.newCDPI, CDPI = st.CDPI;
      st.updateEnd = (newCDPI >= n - 1) ? 64 : (int)CL[newCDPI + 1];
      st.curEntryVal = (int)entry[st.CCL];
      st.numCCL = newCDPI - CDPI + 1;
      CW[newCDPI] = (CDPI == 0) ? 0ull : CW[CDPI];
    }
    __syncthreads();
    { // Operation6
      const int CDPI = st.CDPI, newCDPI = st.newCDPI, CCL = st.CCL;
      const int updateEnd = st.updateEnd;
      const unsigned long long ev =
          (unsigned long long)(st.curEntryVal + st.numCCL);
      const unsigned long long cwn = CW[newCDPI];
      for (int i = CDPI + tid; i < newCDPI; i += CB_THREADS)
        CW[i] = cwn + (unsigned long long)(newCDPI - i);
      if (tid > CCL && tid < updateEnd)
        entry[tid] = ev;
      if (tid == 0 && updateEnd < 64)
        entry[updateEnd] = ev;
    }
    __syncthreads();
    { // Operation7
      const int CCL = st.CCL, CDPI = st.CDPI, updateEnd = st.updateEnd;
      if (tid == CCL)
        first[CCL] = CW[CDPI] ^ ((1ull << CL[CDPI]) - 1);
      if (tid > CCL && tid < updateEnd)
        first[tid] = ~0ull;
    }
    __syncthreads();
    if (tid == 0) { // Operation8
      int newCDPI = st.newCDPI;
      if (newCDPI < n - 1) {
        int CLDiff = (int)CL[newCDPI + 1] - (int)CL[newCDPI];
        CW[newCDPI + 1] = ((CW[st.CDPI] + 1) << CLDiff);
        st.CCL = CL[newCDPI + 1];
        ++newCDPI;
      }
      st.CDPI = newCDPI;
      st.newCDPI = n - 1;
    }
    __syncthreads();
  }
  // Operation9 + Operation10 + ReverseArray(codebook) + ReverseArray(qcode) +
  // ReorderByIndex (GetCodebook.hpp:127-143), collapsed into one gather:
  // after Operation9 CW[i] belongs to the i-th MOST frequent symbol, which is
  // sorted entry dict-1-i.
  for (int t = tid; t < dict; t += CB_THREADS) {
    unsigned sym = A.ssym[dict - 1 - t];
    keys[t] = sym;
    if (t < n) {
      unsigned long long cl = CL[t];
      unsigned long long w =
          (CW[t] | ((cl & 0xffull) << 56)) ^ ((1ull << cl) - 1);
      codebook[sym] = w;
    }
  }
}

size_t mgx_codebook_ws_bytes(int dict, int P) {
  size_t o = 0;
  auto add = [&](size_t b) { o += (b + 15) & ~(size_t)15; };
  add((size_t)dict * 2);
  add((size_t)dict * 4);
  add((size_t)dict);
  add((size_t)dict * 2);
  add((size_t)dict * 2);
  add((size_t)(dict + 2) * 2);
  add((size_t)(2 * (dict / 32 + 2)) * 4);
  add((size_t)P * 8 + 16);
  return o;
}

// host launcher (keeps the kernel symbol private to this translation unit)
cudaError_t mgx_launch_codebook(const uint32_t *d_freq, int dict, int P,
                                uint64_t *d_codebook, uint64_t *d_decodebook,
                                unsigned char *gws, int use_smem, size_t smem,
                                uint64_t *d_scalars, cudaStream_t stream) {
  if (smem) {
    cudaError_t e = cudaFuncSetAttribute(
        k_codebook, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess)
      return e;
  }
  k_codebook<<<1, CB_THREADS, smem, stream>>>(
      d_freq, dict, P, (unsigned long long *)d_codebook,
      (unsigned long long *)d_decodebook, gws, use_smem,
      (unsigned long long *)d_scalars);
  return cudaGetLastError();
}
