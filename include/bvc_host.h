/* include/bvc_host.h -- host side of the `basetype --rerun` drop-in: what sits between the reference's
 * batch files and the GPU calling core (include/bvc_gpu.h), and between the result records and the
 * reference's .cvg / .vcf text.  Implemented in C++ (basevarc_b200/csrc/host), exported as plain C.
 *
 * Reference code each entry point replaces (paths under /root/reference):
 *   bvc_packer_*        the tokeniser / merger of the window loop          src/BaseVarC.cpp:393-441
 *                       (AlleleInfo, AlleleInfoVector, DepM idx            src/BamProcess.h:30-39, src/BaseType.h:21)
 *   bvc_format_cvg      the CVG line of bt_f                               src/BaseVarC.cpp:568-575,609-610,640,662
 *   bvc_format_vcf      the text part of WriteVcf                          src/BaseType.cpp:143-163,198-233
 *   bvc_basetype_main   runBaseType + bt_s for `--rerun` (compute phase)   src/BaseVarC.cpp:176-314,316-465
 *
 * None of these computes a statistic: every number printed comes from a bvc_site_result the GPU produced
 * (or, in tests only, from the oracle).  bvc_basetype_main has no CPU calling path: without a usable
 * sm_100 device it fails with the error of bvc_create.
 */
#ifndef BVC_HOST_H
#define BVC_HOST_H

#include <stddef.h>
#include <stdint.h>
#include "bvc_gpu.h"

#ifdef __cplusplus
extern "C" {
#endif

/* ---- packer: batch-file lines -> structure-of-arrays tile --------------------------------------------------
 * One packer per window (= per reference thread directory): it carries the reference's reused `AlleleInfo ai`
 * across tokens, lines and sites (quirk Q1: an indel token keeps base/mapq/qual/rpr/strand of the last numeric
 * token, src/BaseVarC.cpp:393,432-435).  Before the first numeric token those fields are uninitialised in the
 * reference; here they are zero. */
typedef struct bvc_packer bvc_packer;

bvc_packer* bvc_packer_create(int32_t n_samples);
void bvc_packer_destroy(bvc_packer* pk);
/* Drop the packed sites, keep the stale-field state (call between tiles of one window). */
void bvc_packer_clear(bvc_packer* pk);
/* Start a site: ref_base = BASE_INT8_TABLE[refseq[p - rg_s]] (0..3), pos = 1-based position. */
void bvc_packer_begin_site(bvc_packer* pk, int32_t pos, uint8_t ref_base);
/* Tokens of one batch-file line (no '\n'); sample indices continue across the lines of a site.
 * Returns 0, or BVC_ERR_ARG for a token the GPU path does not accept (non-indel base > 4, qual > 127). */
int  bvc_packer_add_line(bvc_packer* pk, const char* line, size_t len);
/* Close the site.  Returns the number of entries it got (0: the reference prints nothing for it). */
int32_t bvc_packer_end_site(bvc_packer* pk);
int32_t bvc_packer_n_sites(const bvc_packer* pk);
int64_t bvc_packer_n_entries(const bvc_packer* pk);
/* The tile over the packer's arrays (valid until the next packer call) and the host-only side arrays. */
void bvc_packer_tile(const bvc_packer* pk, bvc_tile* out);
const int32_t* bvc_packer_positions(const bvc_packer* pk);             /* [n_sites] */
const char* bvc_packer_indel(const bvc_packer* pk, int64_t entry);     /* indel token text, or NULL */
const char* bvc_packer_error(const bvc_packer* pk);

/* ---- formatter: result record + the site's entries -> the reference's text ---------------------------------
 * Both return the number of bytes the line needs (excluding the NUL); at most cap-1 bytes are written.
 * group_depth / group_af: [n_groups][4] for this site as bvc_fetch_groups returns them, or NULL;
 * group_names: n_groups names in the reference's (sorted) order. */
size_t bvc_format_cvg(char* buf, size_t cap, const char* chr, int32_t pos, const bvc_tile* tile, int32_t site,
                      const bvc_site_result* rec, const bvc_packer* indel_src,
                      int32_t n_groups, const int32_t* group_depth);
/* n_samples = N (all samples get a column, src/BaseType.cpp:152).  Returns 0 when !rec->lrt_ok. */
size_t bvc_format_vcf(char* buf, size_t cap, const char* chr, int32_t pos, const bvc_tile* tile, int32_t site,
                      const bvc_site_result* rec, int32_t n_samples,
                      int32_t n_groups, const char* const* group_names, const int32_t* group_depth,
                      const double* group_af);

/* One number as the reference prints it with "{:.Nf}" (src/BaseType.cpp:208-229, src/BaseVarC.cpp:609): printf's %.Nf,
 * except that the fmt 5.3.1 vendored by the reference prints ONE extra digit for tiny non-zero values
 * (fmt/include/fmt/format-inl.h:591-605, format.h:1092-1105), e.g. 9.6e-5 -> "0.0001" with N = 3. */
size_t bvc_format_fixed(char* buf, size_t cap, double v, int32_t digits);

/* ---- writer of the reference's `--load` batch files from a packed tile (fixtures, benchmarks, CPU-baseline runs of the
 * reference binary): PREFIX.tmp.thread.{t}/batch.{j} as bt_r writes them (src/BaseVarC.cpp:467-534), samples named
 * S0..S{N-1}, indel entries printed as `indel_token` (NULL = "+A").  tile->sample is required. */
int bvc_write_batch_files(const bvc_tile* tile, int32_t n_samples, const char* prefix, int32_t threads, int32_t batch,
                          int32_t level, const char* indel_token);

/* ---- packed tile files: the binary packed batch format ---------------------------------------
 * The reference's `--load` writes one text token per (site, sample) into BGZF batch files (src/BaseVarC.cpp:509-527)
 * and every `--rerun` inflates and tokenises them again (:403-441).  A packed file holds the same information as
 * ready-made tiles in the structure-of-arrays form the GPU library takes, so a rerun reads arrays straight into
 * page-locked buffers.  One file per reference thread directory: PREFIX.tmp.thread.{t}/packed.bvc, written by
 * `BaseVarC_b200 basetype --pack` (from the text batch files, no GPU needed), by bvc_write_packed_files (fixtures), or
 * by a `--load` patched as INTEGRATION.md shows; `--rerun` reads them whenever every window has a complete one
 * (`--no_packed` forces the text files).  Little-endian; every array padded to 16 bytes:
 *   file header (64 B): "BVCPACK1", version = 1, n_samples, n_tiles (written LAST: 0 = incomplete), max_sites,
 *                       max_entries (u64), names_bytes; then the window's sample names, tab-separated
 *   per tile:  header (64 B): "TILE", n_sites S, n_entries R, n_dict, n_indel, indel_bytes, qual dictionary [32]
 *              pos i32[S] | ref_base u8[S] | entry_off u32[S+1] | covered-sample bitmap u8[S][ceil(N/8)]
 *              (entries of a site are in ascending sample order, so N bits replace a 4-byte sample index per entry)
 *              packed8 u8[R] (n_dict > 0; BVC_PACK8) or code u8[R] + qual u8[R] (n_dict == 0: > 30 distinct quals)
 *              mapq u8[R] | rpr u8[R]
 *              indel entries: index u32[n_indel] | stale qual u8[n_indel] | text offsets u32[n_indel+1] | token text
 * 3.0 B per entry + N/8 B per site against ~14 B of text per token. */
typedef struct bvc_packed_reader bvc_packed_reader;
bvc_packed_reader* bvc_packed_open(const char* path, int32_t* n_samples, int32_t* n_tiles, int32_t* max_sites, int64_t* max_entries);
/* Replaces the packer's sites with the next tile: 1, 0 at the end of the file, -1 on a corrupt file (bvc_packed_error). */
int  bvc_packed_next(bvc_packed_reader* r, bvc_packer* into);
const char* bvc_packed_names(const bvc_packed_reader* r);
const char* bvc_packed_error(const bvc_packed_reader* r);
void bvc_packed_close(bvc_packed_reader* r);
int  bvc_packed_is_complete(const char* path, int32_t n_samples);
/* A tile -> PREFIX.tmp.thread.{t}/packed.bvc for t < threads (window partition of src/BaseVarC.cpp:399-403), tiles of
 * tile_sites sites; pos [n_sites] (NULL: 1, 2, ...); samples named S0..S{N-1}; tile->sample, code and qual are required. */
int  bvc_write_packed_files(const bvc_tile* tile, int32_t n_samples, const int32_t* pos, const char* prefix, int32_t threads,
                            int32_t tile_sites, const char* indel_token);

/* ---- BGZF block decoding --------------------------------------------------------------------
 * The reference reads its batch files with htslib's bgzf_getline, i.e. zlib's inflate (src/BaseVarC.cpp:407,
 * SeqLib/htslib/bgzf.c).  The drop-in's reader inflates a block with a whole-buffer DEFLATE decoder written from RFC 1951
 * (csrc/host/fast_inflate.cc: 2.2x zlib on batch-file text) and checks its CRC-32 by carry-less multiplication; a block the
 * decoder declines, or whose CRC does not match, is inflated again by zlib, whose answer stands (BVC_ZLIB_INFLATE=1: zlib only).
 * These two are that decoder and that CRC on caller buffers, for tests:
 * bvc_fast_inflate: raw deflate stream in[0, in_len) -> out[0, *out_len), 1 = decoded, 0 = declined (malformed stream, output
 * larger than out_cap, or a construct it leaves to zlib); bvc_fast_crc32 = zlib's crc32(0, buf, len). */
int bvc_fast_inflate(const uint8_t* in, size_t in_len, uint8_t* out, size_t out_cap, size_t* out_len);
uint32_t bvc_fast_crc32(const uint8_t* buf, size_t len);

/* ---- the command: `BaseVarC basetype --rerun ...` with the compute phase on the GPU(s) ----------------------
 * argv as the reference's runBaseType sees it (argv[0] = "basetype").  Returns the process exit status. */
int bvc_basetype_main(int argc, char** argv);

#ifdef __cplusplus
}
#endif
#endif /* BVC_HOST_H */
