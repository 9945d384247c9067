/*
 * binary_io.cpp — checkpoint I/O in the pll-modules binary format, straight from / into HBM.
 *
 * Replaces src/binary/pll_binary.c (containers and blocks: :60-1290) and
 * src/binary/binary_io_operations.c (field order of a partition block: :161-310, node records:
 * :383-412).  The on-disk layout is the reference's, byte for byte in structure:
 *
 *   file    = pll_binary_header_t | block map (random access only) | blocks...
 *   block   = pll_block_header_t | payload                      (block_len: see each writer)
 *   partition payload = 12 descriptor words | model arrays | [tip codes, charmap, tip-tip lookup,
 *                       tipmap, CLVs, scalers] (DUMP_CLV) | [pattern weights] (DUMP_WGT) | [invariant]
 *   tree payload      = post-order node records (pll_unode_t image, label length, label)
 *
 * What differs is where the bytes come from: CLVs, scale buffers and P-matrices live only on the
 * device, so a dump streams them D2H through one bounce buffer and a load uploads them (and
 * rebuilds the kernels' interleaved P-matrix copies); host-side mirrors (tip codes, weights,
 * invariant sites, eigen data) are marked dirty so the next kernel sees them.
 *
 * Reference quirks kept because they are part of the format: the tip-tip lookup table is written
 * (zeros here: it is scratch that libpll rebuilds before every tip-tip update, core_partials.c:41-103);
 * only the first PLL_ASCII_SIZE BYTES of the 64-bit tipmap are stored (binary_io_operations.c:262);
 * a per-rate scale buffer is stored only up to its first sites_alloc words (:296-298);
 * node records contain raw pointers that only serve as "is inner" / "has label" flags.
 */
#include "internal.h"

#include <cassert>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

namespace
{
enum class Dir
{
  Write,
  Read
};

/* one transfer in the direction of the stream; all sizes in bytes */
struct Stream
{
  FILE * f;
  Dir dir;
  bool ok = true;

  bool io(void * data, size_t bytes)
  {
    if (!ok || !bytes) return ok;
    const size_t n = dir == Dir::Write ? fwrite(data, 1, bytes, f) : fread(data, 1, bytes, f);
    if (n != bytes)
    {
      ok = false;
      pllmod_set_error(PLLMOD_BIN_ERROR_BINARY_IO, "Binary file I/O error while trying to %s data at offset %ld",
                       dir == Dir::Write ? "write" : "read", ftell(f));
    }
    return ok;
  }
  template <typename T>
  bool io(T * data, size_t count)
  {
    return io((void *)data, sizeof(T) * count);
  }
  /* bytes that exist only in the file: zeros out, discarded in */
  bool filler(size_t bytes)
  {
    std::vector<unsigned char> z(bytes, 0);
    return io(z.data(), bytes);
  }
};

bool block_header_io(Stream & s, pll_block_header_t * h)
{
  memcpy(h->pad, "000", 4);
  return s.io(h, 1);
}

/* ++n_blocks in the file header and, when asked, a new (id, offset) row in the block map; the
 * stream position is left where it was (binary_io_operations.c:71-139) */
bool register_block(FILE * f, const pll_block_header_t * bh)
{
  const long here = ftell(f);
  pll_binary_header_t fh;
  if (fseek(f, 0, SEEK_SET) || fread(&fh, sizeof(fh), 1, f) != 1 || fseek(f, 0, SEEK_SET))
  {
    pllmod_set_error(PLLMOD_BIN_ERROR_BINARY_IO, "Cannot read the binary file header");
    return false;
  }
  const unsigned int row = fh.n_blocks++;
  if (fwrite(&fh, sizeof(fh), 1, f) != 1)
  {
    pllmod_set_error(PLLMOD_BIN_ERROR_BINARY_IO, "Cannot update the binary file header");
    return false;
  }
  if (bh && (bh->attributes & PLLMOD_BIN_ATTRIB_UPDATE_MAP))
  {
    if (row >= fh.max_blocks)
    {
      pllmod_set_error(PLLMOD_BIN_ERROR_INVALID_SIZE, "Block map is full (%u blocks)", fh.max_blocks);
      return false;
    }
    pll_block_map_t entry;
    entry.block_id = bh->block_id;
    entry.block_offset = here;
    if (fseek(f, (long)(row * sizeof(pll_block_map_t)), SEEK_CUR) || fwrite(&entry, sizeof(entry), 1, f) != 1)
    {
      pllmod_set_error(PLLMOD_BIN_ERROR_BINARY_IO, "Cannot update the block map");
      return false;
    }
  }
  if (fseek(f, here, SEEK_SET))
  {
    pllmod_set_error(PLLMOD_BIN_ERROR_BINARY_IO, "Cannot return to the block position");
    return false;
  }
  return true;
}

long find_block(FILE * f, int block_id)
{
  unsigned int n = 0;
  pll_block_map_t * map = pllmod_binary_get_map(f, &n);
  if (!map) return PLLMOD_BIN_INVALID_OFFSET;
  long off = PLLMOD_BIN_INVALID_OFFSET;
  for (unsigned int i = 0; i < n && off == PLLMOD_BIN_INVALID_OFFSET; ++i)
    if (map[i].block_id == block_id) off = map[i].block_offset;
  free(map);
  return off;
}

/* position the stream for a load: 0 = here, SEEK = through the map, otherwise an absolute offset */
bool goto_block(FILE * f, int block_id, long offset, int missing_code)
{
  assert(offset >= 0 || offset == PLLMOD_BIN_ACCESS_SEEK);
  if (!offset) return true;
  if (offset == PLLMOD_BIN_ACCESS_SEEK)
  {
    offset = find_block(f, block_id);
    if (offset == PLLMOD_BIN_INVALID_OFFSET)
    {
      pllmod_set_error(missing_code, "Cannot retrieve offset for block %d", block_id);
      return false;
    }
  }
  return fseek(f, offset, SEEK_SET) == 0;
}

bool expect_type(const pll_block_header_t & h, unsigned int type)
{
  if (h.type == type) return true;
  pllmod_set_error(PLLMOD_BIN_ERROR_BLOCK_MISMATCH, "Block type is %d and should be %d", h.type, type);
  return false;
}

size_t clv_doubles(const pll_partition_t * p)
{
  const size_t sites_alloc = p->sites + (p->asc_bias_alloc ? p->states : 0);
  return sites_alloc * p->states_padded * p->rate_cats;
}
size_t scaler_words(const pll_partition_t * p)
{
  const size_t sites_alloc = p->sites + (p->asc_bias_alloc ? p->states : 0);
  return sites_alloc * ((p->attributes & PLL_ATTRIB_RATE_SCALERS) ? p->rate_cats : 1);
}
size_t ttlookup_doubles(const pll_partition_t * p)
{
  const unsigned int l2 = (unsigned int)ceil(log2((double)p->maxstates));
  return ((size_t)1 << (2 * l2)) * p->states_padded * p->rate_cats;
}

/* one device-resident CLV through the stream */
bool clv_io(Stream & s, pll_partition_t * p, unsigned int clv_index, std::vector<double> & bounce)
{
  bounce.resize(clv_doubles(p));
  if (s.dir == Dir::Write)
    return pllb200_download_clv(p, clv_index, bounce.data()) && s.io(bounce.data(), bounce.size());
  return s.io(bounce.data(), bounce.size()) && pllb200_upload_clv(p, clv_index, bounce.data());
}

struct Descriptor
{
  unsigned int tips, clv_buffers, states, sites, rate_matrices, prob_matrices, rate_cats, scale_buffers,
      attributes, states_padded;
  int asc_bias_alloc;
  unsigned int maxstates;
};

/* binary_io_operations.c:161-190 */
bool descriptor_io(Stream & s, Descriptor * d)
{
  return s.io(&d->tips, 1) && s.io(&d->clv_buffers, 1) && s.io(&d->states, 1) && s.io(&d->sites, 1) &&
         s.io(&d->rate_matrices, 1) && s.io(&d->prob_matrices, 1) && s.io(&d->rate_cats, 1) &&
         s.io(&d->scale_buffers, 1) && s.io(&d->attributes, 1) && s.io(&d->states_padded, 1) &&
         s.io(&d->asc_bias_alloc, 1) && s.io(&d->maxstates, 1);
}

Descriptor describe(const pll_partition_t * p)
{
  Descriptor d;
  d.tips = p->tips;
  d.clv_buffers = p->clv_buffers;
  d.states = p->states;
  d.sites = p->sites;
  d.rate_matrices = p->rate_matrices;
  d.prob_matrices = p->prob_matrices;
  d.rate_cats = p->rate_cats;
  d.scale_buffers = p->scale_buffers;
  d.attributes = p->attributes;
  d.states_padded = p->states_padded;
  d.asc_bias_alloc = p->asc_bias_alloc;
  d.maxstates = (p->attributes & PLL_ATTRIB_PATTERN_TIP) ? p->maxstates : 0;
  return d;
}
}  // namespace

namespace
{
/* tables a PATTERN_TIP partition needs before tip codes can be read into it (pll.c:875-909 allocates
 * them on the first pll_set_tip_states) */
bool ensure_tip_tables(pll_partition_t * p)
{
  const size_t sites_alloc = p->sites + (p->asc_bias_alloc ? p->states : 0);
  if (!p->charmap) p->charmap = (unsigned char *)calloc(PLL_ASCII_SIZE, 1);
  if (!p->tipmap) p->tipmap = (pll_state_t *)calloc(PLL_ASCII_SIZE, sizeof(pll_state_t));
  if (!p->tipchars) p->tipchars = (unsigned char **)calloc(p->tips ? p->tips : 1, sizeof(unsigned char *));
  bool ok = p->charmap && p->tipmap && p->tipchars;
  for (unsigned int i = 0; ok && i < p->tips; ++i)
    if (!p->tipchars[i]) ok = (p->tipchars[i] = (unsigned char *)calloc(sites_alloc, 1)) != nullptr;
  if (!ok) pllmod_set_error(PLL_ERROR_MEM_ALLOC, "Cannot allocate space for storing tip characters.");
  return ok;
}

/* everything after the descriptor (binary_io_operations.c:192-310) */
bool body_io(Stream & s, pll_partition_t * p, unsigned int attributes)
{
  PartCtx * c = ctx_of(p);
  const unsigned int S = p->states, sp = p->states_padded, K = p->rate_cats, M = p->rate_matrices;
  const size_t sites_alloc = p->sites + (p->asc_bias_alloc ? p->states : 0);
  const bool writing = s.dir == Dir::Write;

  if (writing)
    for (unsigned int m = 0; m < M; ++m)
      if (!p->eigen_decomp_valid[m] && !pll_update_eigen(p, m)) return false;

  s.io(p->eigen_decomp_valid, M);
  for (unsigned int m = 0; m < M; ++m) s.io(p->eigenvecs[m], (size_t)S * sp);
  for (unsigned int m = 0; m < M; ++m) s.io(p->inv_eigenvecs[m], (size_t)S * sp);
  for (unsigned int m = 0; m < M; ++m) s.io(p->eigenvals[m], sp);
  if (!s.ok) return false;

  /* all P-matrices, one contiguous device array in the reference's layout */
  {
    std::vector<double> pm((size_t)p->prob_matrices * K * S * sp);
    if (writing)
    {
      if (!pllb200_download_pmatrices(p, pm.data()) || !s.io(pm.data(), pm.size())) return false;
    }
    else if (!s.io(pm.data(), pm.size()) || !pllb200_upload_pmatrices(p, pm.data()))
      return false;
  }

  for (unsigned int m = 0; m < M; ++m) s.io(p->subst_params[m], (size_t)S * (S - 1) / 2);
  for (unsigned int m = 0; m < M; ++m) s.io(p->frequencies[m], sp);
  s.io(p->rates, K);
  s.io(p->rate_weights, K);
  s.io(p->prop_invar, M);
  if (!s.ok) return false;
  if (!writing)
    for (unsigned int m = 0; m < M; ++m) ++c->eigen_version[m]; /* the device copy is stale now */

  if (attributes & PLLMOD_BIN_ATTRIB_PARTITION_DUMP_CLV)
  {
    unsigned int first_clv = 0;
    if (p->attributes & PLL_ATTRIB_PATTERN_TIP)
    {
      if (writing && !p->tipchars)
      {
        pllmod_set_error(PLLMOD_BIN_ERROR_LOADSTORE, "Tip states were never set: nothing to dump");
        return false;
      }
      if (!writing && !ensure_tip_tables(p)) return false;
      for (unsigned int t = 0; t < p->tips; ++t) s.io(p->tipchars[t], sites_alloc);
      s.io(p->charmap, PLL_ASCII_SIZE);
      s.filler(ttlookup_doubles(p) * sizeof(double));
      s.io((void *)p->tipmap, PLL_ASCII_SIZE); /* sic: bytes, not entries */
      if (!s.ok) return false;
      first_clv = p->tips;
    }
    std::vector<double> bounce;
    for (unsigned int i = first_clv; i < p->tips + p->clv_buffers; ++i)
      if (!clv_io(s, p, i, bounce)) return false;
    /* sic: the format holds sites_alloc words per scale buffer even with PLL_ATTRIB_RATE_SCALERS, i.e.
     * only the first quarter of a per-rate buffer (binary_io_operations.c:296-298 sizes the transfer
     * with pll_get_sites_number); the rest of the device buffer is left as it is on load */
    std::vector<unsigned int> sc(scaler_words(p));
    for (unsigned int i = 0; i < p->scale_buffers; ++i)
    {
      if (!pllb200_download_scaler(p, i, sc.data())) return false;
      if (!s.io(sc.data(), sites_alloc)) return false;
      if (!writing && !pllb200_upload_scaler(p, i, sc.data())) return false;
    }
  }

  if (attributes & PLLMOD_BIN_ATTRIB_PARTITION_DUMP_WGT) s.io(p->pattern_weights, sites_alloc);

  bool any_pinv = false;
  for (unsigned int m = 0; m < M; ++m) any_pinv |= p->prop_invar[m] > 0;
  if (any_pinv)
  {
    if (!p->invariant)
    {
      if (writing)
      {
        if (!pll_update_invariant_sites(p)) return false;
      }
      else if (!(p->invariant = (int *)malloc(sizeof(int) * p->sites)))
      {
        pllmod_set_error(PLL_ERROR_MEM_ALLOC, "Cannot allocate the invariant site array.");
        return false;
      }
    }
    s.io(p->invariant, p->sites);
  }
  if (!s.ok) return false;
  /* host arrays changed under the device mirrors */
  return writing || pllb200_sync_host_arrays(p) == PLL_SUCCESS;
}

/* node record: struct image, label length, label bytes (binary_io_operations.c:383-412) */
bool node_io(Stream & s, pll_unode_t * node)
{
  unsigned long label_len = 0;
  if (s.dir == Dir::Write)
  {
    if (node->label) label_len = strlen(node->label);
    return s.io(node, 1) && s.io(&label_len, 1) && s.io(node->label, label_len);
  }
  if (!s.io(node, 1) || !s.io(&label_len, 1)) return false;
  /* the stored pointers are only flags; a label of length 0 is no label */
  char * label = nullptr;
  if (label_len)
  {
    label = (char *)malloc(label_len + 1);
    if (!label)
    {
      pllmod_set_error(PLL_ERROR_MEM_ALLOC, "Cannot allocate a node label.");
      return false;
    }
    if (!s.io(label, label_len))
    {
      free(label);
      return false;
    }
    label[label_len] = '\0';
  }
  node->label = label;
  node->data = nullptr;
  return true;
}

int visit_all(pll_unode_t *) { return 1; }

/* shared by clv_dump and pernoderepeats_dump (no site repeats here: both describe a full CLV) */
int dump_clv_block(FILE * f, int block_id, pll_partition_t * p, unsigned int clv_index, unsigned int attributes)
{
  PartCtx * c = ctx_of(p);
  if (clv_index >= p->tips + p->clv_buffers || clv_index < c->clv_first)
  {
    pllmod_set_error(PLLMOD_BIN_ERROR_INVALID_INDEX, "Invalid CLV index");
    return PLL_FAILURE;
  }
  pll_block_header_t bh;
  memset(&bh, 0, sizeof(bh));
  bh.block_id = block_id;
  bh.type = PLLMOD_BIN_BLOCK_CLV;
  bh.attributes = attributes;
  bh.block_len = clv_doubles(p) * sizeof(double);
  Stream s{f, Dir::Write};
  std::vector<double> bounce;
  if (!register_block(f, &bh) || !block_header_io(s, &bh) || !clv_io(s, p, clv_index, bounce))
  {
    if (!pll_errno) pllmod_set_error(PLLMOD_BIN_ERROR_LOADSTORE, "Error loading/storing CLV");
    return PLL_FAILURE;
  }
  return PLL_SUCCESS;
}
}  // namespace

/* ---- containers (pll_binary.c:60-205) ---------------------------------------------------------- */

extern "C" FILE * pllmod_binary_create(const char * filename, pll_binary_header_t * header,
                                       unsigned int access_type, unsigned int n_blocks)
{
  memset(header, 0, sizeof(*header));
  header->access_type = access_type;
  header->max_blocks = n_blocks;
  header->map_offset = access_type == PLLMOD_BIN_ACCESS_RANDOM ? (long)(n_blocks * sizeof(pll_block_map_t)) : 0;
  header->n_blocks = 0;
  if (access_type == PLLMOD_BIN_ACCESS_RANDOM && n_blocks == 0)
  {
    pllmod_set_error(PLLMOD_BIN_ERROR_INVALID_SIZE, "Number of blocks for random access must be greater than 0");
    return nullptr;
  }
  FILE * f = fopen(filename, "w+b");
  if (!f)
  {
    pllmod_set_error(PLL_ERROR_FILE_OPEN, "Cannot open file for writing");
    return nullptr;
  }
  /* reserve the (still empty) block map behind the header */
  if (fwrite(header, sizeof(*header), 1, f) != 1 || fseek(f, header->map_offset, SEEK_CUR))
  {
    pllmod_set_error(PLLMOD_BIN_ERROR_BINARY_IO, "Error writing header to file");
    fclose(f);
    return nullptr;
  }
  return f;
}

extern "C" FILE * pllmod_binary_open(const char * filename, pll_binary_header_t * header)
{
  FILE * f = fopen(filename, "rb");
  if (!f)
  {
    pllmod_set_error(PLL_ERROR_FILE_OPEN, "Cannot open file for reading");
    return nullptr;
  }
  if (fread(header, sizeof(*header), 1, f) != 1)
  {
    pllmod_set_error(PLLMOD_BIN_ERROR_BINARY_IO, "Error reading header from file");
    fclose(f);
    return nullptr;
  }
  fseek(f, header->map_offset, SEEK_CUR);
  return f;
}

extern "C" FILE * pllmod_binary_append_open(const char * filename, pll_binary_header_t * header)
{
  FILE * f = fopen(filename, "r+b");
  if (!f)
  {
    pllmod_set_error(PLL_ERROR_FILE_OPEN, "Cannot open file for appending");
    return nullptr;
  }
  if (fread(header, sizeof(*header), 1, f) != 1)
  {
    pllmod_set_error(PLLMOD_BIN_ERROR_BINARY_IO, "Error reading header from file");
    fclose(f);
    return nullptr;
  }
  /* behind the last block, but never inside the reserved map (pll_binary.c:172-186) */
  long pos = fseek(f, 0, SEEK_END) ? -1 : ftell(f);
  if (pos >= 0 && header->map_offset > pos) pos = header->map_offset;
  if (pos < 0 || fseek(f, pos, SEEK_SET))
  {
    pllmod_set_error(PLLMOD_BIN_ERROR_BINARY_IO, "Cannot position the file for appending");
    fclose(f);
    return nullptr;
  }
  return f;
}

extern "C" int pllmod_binary_close(FILE * bin_file) { return fclose(bin_file); }

extern "C" pll_block_map_t * pllmod_binary_get_map(FILE * bin_file, unsigned int * n_blocks)
{
  pll_binary_header_t fh;
  if (fseek(bin_file, 0, SEEK_SET) || fread(&fh, sizeof(fh), 1, bin_file) != 1) return nullptr;
  pll_block_map_t * map = (pll_block_map_t *)malloc(sizeof(pll_block_map_t) * (fh.n_blocks ? fh.n_blocks : 1));
  if (!map) return nullptr;
  if (fh.n_blocks && fread(map, sizeof(pll_block_map_t), fh.n_blocks, bin_file) != fh.n_blocks)
  {
    free(map);
    return nullptr;
  }
  *n_blocks = fh.n_blocks;
  return map;
}

/* ---- partition blocks (pll_binary.c:207-505) ---------------------------------------------------- */

extern "C" int pllmod_binary_partition_dump(FILE * bin_file, int block_id, pll_partition_t * partition,
                                            unsigned int attributes)
{
  if (!partition || ctx_of(partition)->magic != PLLB200_MAGIC)
  {
    pllmod_set_error(PLL_ERROR_PARAM_INVALID, "Not a libpll_b200 partition.");
    return PLL_FAILURE;
  }
  const long start = ftell(bin_file);
  pll_block_header_t bh;
  memset(&bh, 0, sizeof(bh));
  bh.block_id = block_id;
  bh.type = PLLMOD_BIN_BLOCK_PARTITION;
  bh.attributes = attributes;
  Stream s{bin_file, Dir::Write};
  Descriptor d = describe(partition);
  /* the length is only known afterwards: header, payload, then the header once more */
  if (!register_block(bin_file, &bh) || !block_header_io(s, &bh) || !descriptor_io(s, &d) ||
      !body_io(s, partition, attributes))
    return PLL_FAILURE;
  const long end = ftell(bin_file);
  bh.block_len = (size_t)(end - start);
  if (fseek(bin_file, start, SEEK_SET) || !block_header_io(s, &bh) || fseek(bin_file, end, SEEK_SET))
  {
    if (!pll_errno) pllmod_set_error(PLLMOD_BIN_ERROR_BINARY_IO, "Cannot finalise the partition block header");
    return PLL_FAILURE;
  }
  return PLL_SUCCESS;
}

extern "C" pll_partition_t * pllmod_binary_partition_load(FILE * bin_file, int block_id,
                                                          pll_partition_t * partition,
                                                          unsigned int * attributes, long int offset)
{
  if (*attributes & PLLMOD_BIN_ATTRIB_PARTITION_LOAD_SKELETON)
  {
    /* a skeleton owns no CLV storage; here all CLVs of a partition are one HBM arena */
    pllmod_set_error(PLLMOD_ERROR_NOT_IMPLEMENTED, "PLLMOD_BIN_ATTRIB_PARTITION_LOAD_SKELETON is not supported");
    return nullptr;
  }
  if (!goto_block(bin_file, block_id, offset, PLLMOD_BIN_ERROR_BINARY_IO)) return nullptr;
  Stream s{bin_file, Dir::Read};
  pll_block_header_t bh;
  if (!block_header_io(s, &bh) || !expect_type(bh, PLLMOD_BIN_BLOCK_PARTITION)) return nullptr;
  *attributes = bh.attributes;

  /* the reference reads the descriptor only when it has to create the partition (pll_binary.c:314-330):
   * a caller-provided partition is trusted to match, and the descriptor bytes are then consumed by
   * the body reader there.  That is a latent bug (the body would start 48 bytes early); here the
   * descriptor is always read and checked. */
  Descriptor d;
  if (!descriptor_io(s, &d)) return nullptr;
  pll_partition_t * p = partition;
  if (!p)
  {
    p = pll_partition_create(d.tips, d.clv_buffers, d.states, d.sites, d.rate_matrices, d.prob_matrices,
                             d.rate_cats, d.scale_buffers, d.attributes);
    if (!p) return nullptr;
  }
  const Descriptor have = describe(p);
  if (have.tips != d.tips || have.clv_buffers != d.clv_buffers || have.states != d.states ||
      have.sites != d.sites || have.rate_matrices != d.rate_matrices || have.prob_matrices != d.prob_matrices ||
      have.rate_cats != d.rate_cats || have.scale_buffers != d.scale_buffers ||
      have.states_padded != d.states_padded || (have.asc_bias_alloc != 0) != (d.asc_bias_alloc != 0) ||
      ((have.attributes ^ d.attributes) & (PLL_ATTRIB_PATTERN_TIP | PLL_ATTRIB_RATE_SCALERS)))
  {
    pllmod_set_error(PLLMOD_BIN_ERROR_INVALID_SIZE,
                     "Partition block does not fit: %u tips, %u states (padded %u), %u sites, %u categories in the file",
                     d.tips, d.states, d.states_padded, d.sites, d.rate_cats);
    if (!partition) pll_partition_destroy(p);
    return nullptr;
  }
  p->maxstates = d.maxstates;
  if (!body_io(s, p, *attributes))
  {
    if (!partition) pll_partition_destroy(p);
    return nullptr;
  }
  return p;
}

/* ---- site-repeat blocks: this library rejects PLL_ATTRIB_SITE_REPEATS at partition creation, and
 * for partitions without it the reference's functions succeed without touching the file
 * (pll_binary.c:521-527, 571-576) ------------------------------------------------------------- */

extern "C" int pllmod_binary_repeats_dump(FILE *, int, pll_partition_t *, unsigned int) { return PLL_SUCCESS; }
extern "C" int pllmod_binary_repeats_load(FILE *, int, pll_partition_t *, unsigned int *, long int)
{
  return PLL_SUCCESS;
}
extern "C" int pllmod_binary_pernoderepeats_dump(FILE * bin_file, int block_id, pll_partition_t * partition,
                                                 unsigned int clv_index, unsigned int attributes)
{
  return dump_clv_block(bin_file, block_id, partition, clv_index, attributes);
}
/* pll_binary.c:698-707: a stub that returns 0 there, too */
extern "C" int pllmod_binary_pernoderepeats_load(FILE *, int, pll_partition_t *, unsigned int, unsigned int *,
                                                 long int)
{
  return 0;
}

/* ---- CLV blocks (pll_binary.c:719-865) ----------------------------------------------------------- */

extern "C" int pllmod_binary_clv_dump(FILE * bin_file, int block_id, pll_partition_t * partition,
                                      unsigned int clv_index, unsigned int attributes)
{
  return dump_clv_block(bin_file, block_id, partition, clv_index, attributes);
}

extern "C" int pllmod_binary_clv_load(FILE * bin_file, int block_id, pll_partition_t * partition,
                                      unsigned int clv_index, unsigned int * attributes, long int offset)
{
  assert(partition);
  PartCtx * c = ctx_of(partition);
  if (clv_index >= partition->tips + partition->clv_buffers || clv_index < c->clv_first)
  {
    pllmod_set_error(PLLMOD_BIN_ERROR_INVALID_INDEX, "Invalid CLV index");
    return PLL_FAILURE;
  }
  if (!goto_block(bin_file, block_id, offset, PLLMOD_BIN_ERROR_MISSING_BLOCK)) return PLL_FAILURE;
  Stream s{bin_file, Dir::Read};
  pll_block_header_t bh;
  if (!block_header_io(s, &bh) || !expect_type(bh, PLLMOD_BIN_BLOCK_CLV)) return PLL_FAILURE;
  if (bh.block_len != clv_doubles(partition) * sizeof(double))
  {
    pllmod_set_error(PLLMOD_BIN_ERROR_BLOCK_LENGTH, "Wrong block length");
    return PLL_FAILURE;
  }
  *attributes = bh.attributes;
  std::vector<double> bounce;
  if (!clv_io(s, partition, clv_index, bounce))
  {
    if (!pll_errno) pllmod_set_error(PLLMOD_BIN_ERROR_LOADSTORE, "Error loading/storing CLV");
    return PLL_FAILURE;
  }
  return PLL_SUCCESS;
}

/* ---- tree blocks (pll_binary.c:867-1120) --------------------------------------------------------- */

extern "C" int pllmod_binary_utree_dump(FILE * bin_file, int block_id, pll_unode_t * tree,
                                        unsigned int tip_count, unsigned int attributes)
{
  pll_errno = 0;
  if (!tree->next)
  {
    pllmod_set_error(PLLMOD_BIN_ERROR_BINARY_IO, "Tree should not be a tip node");
    return PLL_FAILURE;
  }
  const unsigned int inner = tip_count - 2, nodes = tip_count + inner;
  std::vector<pll_unode_t *> order(nodes);
  unsigned int visited = 0;
  if (!pll_utree_traverse(tree, PLL_TREE_TRAVERSE_POSTORDER, visit_all, order.data(), &visited) ||
      visited != nodes)
  {
    pllmod_set_error(PLLMOD_BIN_ERROR_BINARY_IO, "Error traversing utree");
    return PLL_FAILURE;
  }
  pll_block_header_t bh;
  memset(&bh, 0, sizeof(bh));
  bh.block_id = block_id;
  bh.type = PLLMOD_BIN_BLOCK_TREE;
  bh.attributes = attributes;
  bh.block_len = (size_t)(tip_count + 3 * inner) * sizeof(pll_unode_t); /* labels are not counted */
  Stream s{bin_file, Dir::Write};
  if (!register_block(bin_file, &bh) || !block_header_io(s, &bh)) return PLL_FAILURE;
  /* tips: one record; inner nodes: the three records of the ring, entry node first */
  for (pll_unode_t * n : order)
  {
    if (!node_io(s, n)) return PLL_FAILURE;
    if (n->next && (!node_io(s, n->next) || !node_io(s, n->next->next))) return PLL_FAILURE;
  }
  return PLL_SUCCESS;
}

extern "C" pll_unode_t * pllmod_binary_utree_load(FILE * bin_file, int block_id, unsigned int * attributes,
                                                  long int offset)
{
  if (!goto_block(bin_file, block_id, offset, PLLMOD_BIN_ERROR_BINARY_IO)) return nullptr;
  Stream s{bin_file, Dir::Read};
  pll_block_header_t bh;
  if (!block_header_io(s, &bh) || !expect_type(bh, PLLMOD_BIN_BLOCK_TREE)) return nullptr;
  const size_t records = bh.block_len / sizeof(pll_unode_t);
  if (records % 4 != 2)
  {
    pllmod_set_error(PLLMOD_BIN_ERROR_BLOCK_LENGTH, "Wrong block length");
    return nullptr;
  }
  const unsigned int tips = (unsigned int)((records + 6) / 4), nodes = 2 * tips - 2;
  *attributes = bh.attributes;

  /* post-order rebuild: tips are pushed, an inner node pops its two children */
  std::vector<pll_unode_t *> stack, made;
  std::vector<char *> labels;
  auto fail = [&]() -> pll_unode_t * {
    for (char * l : labels) free(l);
    for (pll_unode_t * n : made) free(n);
    if (!pll_errno) pllmod_set_error(PLLMOD_BIN_ERROR_BINARY_IO, "Malformed tree block");
    return nullptr;
  };
  auto fresh = [&]() -> pll_unode_t * {
    pll_unode_t * n = (pll_unode_t *)calloc(1, sizeof(pll_unode_t));
    if (n) made.push_back(n);
    return n;
  };
  for (unsigned int i = 0; i < nodes; ++i)
  {
    pll_unode_t * a = fresh();
    if (!a || !node_io(s, a)) return fail();
    if (a->label) labels.push_back(a->label);
    const bool is_inner = a->next != nullptr;
    a->next = a->back = nullptr;
    if (is_inner)
    {
      /* the three records of an inner node share one label: keep the first copy */
      pll_unode_t * ring[2] = {fresh(), fresh()};
      for (pll_unode_t * r : ring)
      {
        if (!r || !node_io(s, r)) return fail();
        free(r->label);
        r->label = a->label;
        r->next = r->back = nullptr;
      }
      if (stack.size() < 2) return fail();
      pll_unode_t *b = ring[0], *c = ring[1];
      a->next = b;
      b->next = c;
      c->next = a;
      pll_unode_t * right = stack.back();
      stack.pop_back();
      pll_unode_t * left = stack.back();
      stack.pop_back();
      c->back = right;
      right->back = c;
      b->back = left;
      left->back = b;
    }
    stack.push_back(a);
  }
  if (stack.size() != 2) return fail();
  pll_unode_t * root = stack[1];
  root->back = stack[0];
  stack[0]->back = root;
  return root;
}

/* ---- custom blocks (pll_binary.c:1122-1263) ------------------------------------------------------- */

extern "C" int pllmod_binary_custom_dump(FILE * bin_file, int block_id, void * data, size_t size,
                                         unsigned int attributes)
{
  pll_block_header_t bh;
  memset(&bh, 0, sizeof(bh));
  bh.block_id = block_id;
  bh.type = PLLMOD_BIN_BLOCK_CUSTOM;
  bh.attributes = attributes;
  bh.block_len = size;
  Stream s{bin_file, Dir::Write};
  return register_block(bin_file, &bh) && block_header_io(s, &bh) && s.io(data, size) ? PLL_SUCCESS
                                                                                       : PLL_FAILURE;
}

extern "C" void * pllmod_binary_custom_load(FILE * bin_file, int block_id, size_t * size, unsigned int * type,
                                            unsigned int * attributes, long int offset)
{
  if (!goto_block(bin_file, block_id, offset, PLLMOD_BIN_ERROR_BINARY_IO)) return nullptr;
  Stream s{bin_file, Dir::Read};
  pll_block_header_t bh;
  if (!block_header_io(s, &bh)) return nullptr;
  *type = bh.type;
  *size = bh.block_len;
  *attributes = bh.attributes;
  void * data = (bh.attributes & PLLMOD_BIN_ATTRIB_ALIGNED) ? pll_aligned_alloc(bh.block_len, PLL_ALIGNMENT_AVX)
                                                            : malloc(bh.block_len ? bh.block_len : 1);
  if (!data)
  {
    pllmod_set_error(PLL_ERROR_MEM_ALLOC, "Cannot allocate space for storing data.");
    return nullptr;
  }
  if (!s.io(data, bh.block_len))
  {
    if (bh.attributes & PLLMOD_BIN_ATTRIB_ALIGNED)
      pll_aligned_free(data);
    else
      free(data);
    return nullptr;
  }
  return data;
}
