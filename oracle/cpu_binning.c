/* TEST INFRASTRUCTURE -- see cpu_raster.h.  Tile binning of the reference
 * (K3 duplicateWithKeys, K4 stable radix sort, K5 identifyTileRanges). */
#include "cpu_raster.h"
#include "cpu_math.h"
#include <stdlib.h>

/* Stable LSD radix sort of (key,value) pairs on the low `nbits` key bits:
 * the result cub::DeviceRadixSort::SortPairs is specified to give
 * (DGR rasterizer_impl.cu:306-311). */
static void radix_sort_pairs(int64_t n, const uint64_t* kin, const uint32_t* vin, uint64_t* kout,
                             uint32_t* vout, int nbits)
{
  uint64_t* ka = (uint64_t*)malloc(sizeof(uint64_t) * (size_t)(n ? n : 1));
  uint32_t* va = (uint32_t*)malloc(sizeof(uint32_t) * (size_t)(n ? n : 1));
  size_t* cnt = (size_t*)malloc(sizeof(size_t) * 65537);
  const uint64_t* ks = kin;
  const uint32_t* vs = vin;
  int npass = (nbits + 15) / 16;
  if (npass == 0) npass = 1;
  for (int pass = 0; pass < npass; pass++) {
    /* ping-pong so that the last pass lands in kout/vout */
    int last = (pass == npass - 1);
    int to_out = ((npass - 1 - pass) % 2) == 0;
    uint64_t* kd = to_out ? kout : ka;
    uint32_t* vd = to_out ? vout : va;
    (void)last;
    int shift = 16 * pass;
    int width = nbits - shift < 16 ? nbits - shift : 16;
    uint64_t mask = width > 0 ? ((1ull << width) - 1) : 0;
    memset(cnt, 0, sizeof(size_t) * 65537);
    for (int64_t i = 0; i < n; i++) cnt[((ks[i] >> shift) & mask) + 1]++;
    for (int d = 0; d < 65536; d++) cnt[d + 1] += cnt[d];
    for (int64_t i = 0; i < n; i++) {
      size_t dst = cnt[(ks[i] >> shift) & mask]++;
      kd[dst] = ks[i];
      vd[dst] = vs[i];
    }
    ks = kd;
    vs = vd;
  }
  free(ka);
  free(va);
  free(cnt);
}

/* DGR rasterizer_impl.cu:35-50 getHigherMsb */
static uint32_t higher_msb(uint32_t n)
{
  uint32_t msb = sizeof(n) * 4, step = msb;
  while (step > 1) {
    step /= 2;
    if (n >> msb) msb += step;
    else msb -= step;
  }
  if (n >> msb) msb++;
  return msb;
}

void orc_bin(const orc_scene* s, const orc_geom* g, int64_t num_rendered, uint64_t* keys_unsorted,
             uint32_t* vals_unsorted, uint64_t* keys_sorted, uint32_t* point_list, uint32_t* ranges)
{
  const int gx = (s->W + ORC_BLOCK_X - 1) / ORC_BLOCK_X;
  const int gy = (s->H + ORC_BLOCK_Y - 1) / ORC_BLOCK_Y;
  /* K3, DGR rasterizer_impl.cu:70-111: key = tile<<32 | float_bits(depth),
   * emitted at offsets[idx-1].. in row-major tile order. */
  for (int idx = 0; idx < s->P; idx++) {
    if (g->radii[idx] <= 0) continue;
    uint32_t off = idx == 0 ? 0 : g->point_offsets[idx - 1];
    int rmin[2], rmax[2];
    get_rect(g->means2D[2 * idx], g->means2D[2 * idx + 1], g->radii[idx], gx, gy, rmin, rmax);
    for (int y = rmin[1]; y < rmax[1]; y++)
      for (int x = rmin[0]; x < rmax[0]; x++) {
        uint64_t key = (uint64_t)(y * gx + x);
        key <<= 32;
        key |= f2bits(g->depths[idx]);
        keys_unsorted[off] = key;
        vals_unsorted[off] = (uint32_t)idx;
        off++;
      }
  }
  /* K4, DGR rasterizer_impl.cu:303-311 */
  int bit = (int)higher_msb((uint32_t)(gx * gy));
  radix_sort_pairs(num_rendered, keys_unsorted, vals_unsorted, keys_sorted, point_list, 32 + bit);
  /* memset + K5, DGR rasterizer_impl.cu:313, 116-138 */
  memset(ranges, 0, sizeof(uint32_t) * 2 * (size_t)gx * gy);
  for (int64_t i = 0; i < num_rendered; i++) {
    uint32_t cur = (uint32_t)(keys_sorted[i] >> 32);
    if (i == 0) ranges[2 * cur] = 0;
    else {
      uint32_t prev = (uint32_t)(keys_sorted[i - 1] >> 32);
      if (cur != prev) {
        ranges[2 * prev + 1] = (uint32_t)i;
        ranges[2 * cur] = (uint32_t)i;
      }
    }
    if (i == num_rendered - 1) ranges[2 * cur + 1] = (uint32_t)num_rendered;
  }
}
