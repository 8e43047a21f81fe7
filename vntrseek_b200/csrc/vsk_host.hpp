// Host-side structures of libvsk: TR sets as psearch holds them after loading,
// the reference seed index, and the clustering / output stage.
#pragma once
#include <stdint.h>

#include <string>
#include <unordered_map>
#include <vector>

#include "../../include/vsk.h"

// One .leb36 file in memory.  Record i owns two profiles: 2i (as written) and
// 2i+1 (its reverse complement, also as written) -- ReadProfileWithRC,
// reference src/psearch1.91/profile.h:448-732.
struct vsk_trset {
    int32_t n = 0;
    bool is_read = false;
    std::vector<int32_t> key, patlen;
    std::vector<int64_t> prof_off;  // [2n+1]
    std::vector<uint16_t> ids;
    std::vector<int64_t> cons_off;  // [2n+1] consensus without dash columns (profile.h:280-322)
    std::vector<char> cons;
    // flanks per profile p: left = fl[fl_off[2p] .. fl_off[2p+1]), right = fl[fl_off[2p+1] .. fl_off[2p+2]).
    // forward: left = reverse(file left flank), right = file right flank;
    // rc: left = complement(forward right), right = complement(forward left) (profile.h:687-695)
    std::vector<int64_t> fl_off;  // [4n+1]
    std::vector<char> fl;
    std::vector<int32_t> leftlen, rightlen;  // [2n] lengths after the MAXFLANKCONSIDERED / REFLEN cut
    // redundancy (rotindex) state, LoadRotated psearch.c:1086-1201
    std::vector<uint8_t> master, dir;          // [n]
    std::vector<std::vector<int32_t>> rotlist; // [n] member record indices, self first
    std::unordered_map<int32_t, int32_t> by_key;
    std::string err;
    // rotlists flattened for the device (built on demand, dropped when the lists change)
    std::vector<int32_t> rot_off, rot_mem;
    bool rot_flat = false;
    // first 64-column block of each consensus (what the LCS kernels index by), built with the flat rotlists
    std::vector<int64_t> cons_blk;  // [2n+1]
    bool blk_ready = false;
    std::vector<void *> pinned;  // host ranges registered with cudaHostRegister (vsk_trset_pin)
    // Raw mode (vsk_trset_load_leb36_raw): profiles and flanks are still the file's text and are decoded on
    // the device when the set is uploaded (LEB36 digits -> ids, consensus, flank clean-up).  ids / cons / fl
    // stay empty on the host; prof_off, fl_off, leftlen, rightlen are known from the field lengths;
    // cons_off is filled in by the first upload (raw_decoded).
    bool raw = false, raw_decoded = false;
    std::vector<char> raw_text;        // the whole file
    std::vector<int64_t> raw_prof;     // [2n] offset of each profile's LEB36 field
    std::vector<int64_t> raw_flank;    // [2n] offset of the left / right flank text of each record
    std::vector<int32_t> raw_flen;     // [2n] their lengths
};

// builds the derived host arrays an upload needs (flat rotlists, cons_blk) if they are not there yet
void vsk_trset_prepare(vsk_trset *ts);

// Seed geometry per TRANGE (psearch.c:161-318)
struct vsk_seedinfo {
    int tsize[15];             // seed length
    int nones[15];             // number of '1's
    int off[15][32];           // foffset[0][k]: non-positive offsets of the '1' positions from the window end
    int16_t trange_of[5001];   // pattern length -> TRANGE (0 = not searched)
    int16_t err_of[5001];      // (int)(pmin*0.10 + 1.0) in double (hclust.h:382-392)
};
const vsk_seedinfo &vsk_seeds();

// Reference seed index flattened for the device.  Buckets hold the DISTINCT
// (pmin, ref profile) entries sorted by (pmin, profile) with the number of seed
// hits each stands for (`mult`), plus, per entry, the profile of the first hit of
// its pmin group in the reference's own bucket order (`first`): the reference's
// window takes one hit beyond `hi`, namely that one (hclust.h:388-421).
struct vsk_refindex_host {
    std::vector<uint64_t> codes[15];      // sorted unique (seedcode << 32 | seedcode2)
    std::vector<int64_t> bucket_off[15];  // [ncodes+1] into the entry arrays (global offsets)
    std::vector<int16_t> e_pmin;
    std::vector<int32_t> e_prof, e_mult, e_first;
    std::vector<int64_t> e_cum;  // exclusive prefix of mult inside each bucket
    int64_t total_hits = 0;
    uint64_t tuples = 0;
};

int vsk_build_refindex_host(const vsk_trset &refs, vsk_refindex_host &out, std::string &err);
// the same through the on-disk cache (vsk_index.cu, "index cache"); cache_path may be null
int vsk_refindex_get(const vsk_trset &refs, const char *cache_path, vsk_refindex_host &hi, std::string &err,
                     int *from_cache);

// accepted link, in the order psearch would have called ClusterBaseAddLink
struct vsk_link_rec {
    int32_t refkey, readkey;
    int32_t dir, lerr, rerr;
    int32_t dsum, sumw;
};
