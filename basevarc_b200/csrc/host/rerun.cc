// rerun.cc -- `BaseVarC basetype --rerun` with the compute phase on the GPU(s): bvc_basetype_main.
//
// What it keeps from the reference's runBaseType / bt_s (paths under /root/reference, src/BaseVarC.cpp):
//   the command line (:129-148,797-827), N = lines of the BAM list (:186-189), the region arithmetic and the
//   site list pv (:190-204), the batch-file names and their BGZF+EOF validity probe (:214-233), the window
//   partition (:399-403), the headers (:65-88,364-391), the status lines on stdout/stderr, --keep_tmp (:458-462),
//   the order of the output (windows in thread order, header from window 0 only, :268-295).
// What it replaces: one std::thread per window calling bt_f per site (:263-266,403-454) becomes one reader /
// tokeniser / formatter thread per window feeding tiles of sites to one dispatcher thread per GPU through the
// C ABI of include/bvc_gpu.h (windows [g*T/G, (g+1)*T/G) -> GPU g, no collective); window outputs are
// compressed in memory as independent BGZF blocks and concatenated, instead of re-read and re-compressed.
// The load phase (BAM decoding, bt_r) is NOT here: it stays with the reference binary (`--load`); this
// command needs complete batch files and says so otherwise.
#include <getopt.h>
#include <sched.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>
#include <unistd.h>

#include <algorithm>
#include <atomic>
#include <chrono>
#include <condition_variable>
#include <deque>
#include <fstream>
#include <future>
#include <iostream>
#include <map>
#include <memory>
#include <mutex>
#include <sstream>
#include <string>
#include <thread>
#include <unordered_map>
#include <vector>

#include "../../../include/bvc_host.h"
#include "bgzf_io.hpp"
#include "fasta.hpp"
#include "format.hpp"
#include "packed_io.hpp"
#include "packer.hpp"

namespace bvc_host {
namespace {

const char* kUsage =
    "Commands: BaseVarC basetype (B200 calling core; needs batch files written by `BaseVarC basetype --load`)\n"
    "Usage   : BaseVarC basetype --rerun [options]\n\n"
    "Options :\n"
    "  --input,      -i         BAM/CRAM file list, one file per row (only its line count N is used)\n"
    "  --output,     -o         Output file prefix\n"
    "  --reference,  -r         Reference file\n"
    "  --region,     -s         Samtools-like region <chr:start-end>\n"
    "  --group,      -g         Population group information <SampleID Group>\n"
    "  --mapq,       -q <INT>   (load phase only)\n"
    "  --thread,     -t <INT>   Number of windows the batch files were written with\n"
    "  --batch,      -b <INT>   Number of samples each batch\n"
    "  --maf,        -a <FLOAT> Minimum allele count frequency [min(0.001, 100/N, maf)]\n"
    "  --rerun,                 Read previous loaded data and rerun\n"
    "  --keep_tmp,              Don't remove tmp files when basetype finished\n"
    "  --gpus <INT>             GPUs to use [all visible; env BVC_GPUS]\n"
    "  --tile_sites <INT>       Sites per GPU tile [auto]\n"
    "  --pack                   Convert the text batch files into packed tile files (PREFIX.tmp.thread.{t}/packed.bvc; no GPU\n"
    "                           needed); --rerun reads them instead of the text files whenever every window has one\n"
    "  --no_packed              --rerun from the text batch files even if packed tile files exist\n"
    "  --pending_mb <INT>       Host memory (MiB, all windows together) for tiles packed ahead while the GPU starts up [4096]\n"
    "  --shard_samples <INT>    Sample-sharded calling for very wide cohorts: rank r of W reads the batch files [r*J/W, (r+1)*J/W)\n"
    "                           of every window, one GPU per rank with NCCL (fewer GPUs than ranks: all ranks on GPU 0)\n";

const char* kCvgHeader =
    "##fileformat=CVGv1.0\n"
    "##Group information is the depth of A:C:G:T:Indel\n"
    "#CHROM\tPOS\tREF\tDepth\tA\tC\tG\tT\tIndels\tFS\tSOR\tStrand_Coverage(REF_FWD,REF_REV,ALT_FWD,ALT_REV)";

const char* kVcfHeader =
    "##fileformat=VCFv4.2\n"
    "##FILTER=<ID=LowQual,Description=\"Low quality (QUAL < 60)\">\n"
    "##FORMAT=<ID=GT,Number=1,Type=String,Description=\"Genotype\">\n"
    "##FORMAT=<ID=AB,Number=1,Type=String,Description=\"Allele Base\">\n"
    "##FORMAT=<ID=SO,Number=1,Type=String,Description=\"Strand orientation of the mapping base. Marked as + or -\">\n"
    "##FORMAT=<ID=BP,Number=1,Type=String,Description=\"Base Probability which calculate by base quality\">\n"
    "##INFO=<ID=CM_AF,Number=A,Type=Float,Description=\"An ordered, comma delimited list of allele frequencies base on LRT algorithm\">\n"
    "##INFO=<ID=CM_CAF,Number=A,Type=Float,Description=\"An ordered, comma delimited list of allele frequencies just base on read count\">\n"
    "##INFO=<ID=CM_AC,Number=A,Type=Integer,Description=\"An ordered, comma delimited allele depth in CMDB\">\n"
    "##INFO=<ID=CM_DP,Number=A,Type=Integer,Description=\"Total Depth\">\n"
    "##INFO=<ID=SB_REF,Number=A,Type=Integer,Description=\"Read number support REF: Forward,Reverse\">\n"
    "##INFO=<ID=SB_ALT,Number=A,Type=Integer,Description=\"Read number support ALT: Forward,Reverse\">\n"
    "##INFO=<ID=FS,Number=1,Type=Float,Description=\"Phred-scaled p-value using Fisher's exact test to detect strand bias\">\n"
    "##INFO=<ID=BaseQRankSum,Number=1,Type=Float,Description=\"Phred-score from Wilcoxon rank sum test of Alt Vs. Ref base qualities\">\n"
    "##INFO=<ID=SOR,Number=1,Type=Float,Description=\"Symmetric Odds Ratio of 2x2 contingency table to detect strand bias\">\n"
    "##INFO=<ID=MQRankSum,Number=1,Type=Float,Description=\"Phred-score From Wilcoxon rank sum test of Alt vs. Ref read mapping qualities\">\n"
    "##INFO=<ID=ReadPosRankSum,Number=1,Type=Float,Description=\"Phred-score from Wilcoxon rank sum test of Alt vs. Ref read position bias\">\n"
    "##INFO=<ID=QD,Number=1,Type=Float,Description=\"Variant Confidence Quality by Depth\">\n";

struct Options {
    bool rerun = false, load = false, keep_tmp = false, verbose = false;
    int shard_samples = 0;                     // --shard_samples W: the sample-sharded compute phase over W ranks (shard_rerun.inc)
    bool pack = false, no_packed = false;      // --pack: text batch files -> packed tile files (no GPU); --no_packed: ignore packed files
    int mapq = 10, thread = 1, batch = 10, gpus = 0, tile_sites = 0;
    int pending_mb = 4096;                     // --pending_mb: host memory all windows together may hold in tiles waiting for the GPU
    double maf = 0.001;
    std::string input, reference, group, region, output;
};

// src/BaseVarC.cpp:797-827 (same letters, same long names; 6/7/8 are the reference's codes for the flags)
bool parse_options(int argc, char** argv, Options& o)
{
    static const option longopts[] = {
        {"verbose", no_argument, nullptr, 'v'},       {"input", required_argument, nullptr, 'i'},
        {"output", required_argument, nullptr, 'o'},  {"reference", required_argument, nullptr, 'r'},
        {"posfile", required_argument, nullptr, 'p'}, {"region", required_argument, nullptr, 's'},
        {"group", required_argument, nullptr, 'g'},   {"mapq", required_argument, nullptr, 'q'},
        {"batch", required_argument, nullptr, 'b'},   {"thread", required_argument, nullptr, 't'},
        {"maf", required_argument, nullptr, 'a'},     {"rerun", no_argument, nullptr, 8},
        {"load", no_argument, nullptr, 7},            {"keep_tmp", no_argument, nullptr, 6},
        {"gpus", required_argument, nullptr, 1},      {"tile_sites", required_argument, nullptr, 2},
        {"pack", no_argument, nullptr, 3},            {"no_packed", no_argument, nullptr, 4},
        {"shard_samples", required_argument, nullptr, 5}, {"pending_mb", required_argument, nullptr, 9},
        {nullptr, 0, nullptr, 0}};
    bool die = false;
    optind = 1;
    for (int c; (c = getopt_long(argc, argv, "hva:i:r:p:s:o:q:t:b:g:", longopts, nullptr)) != -1;) {
        const char* a = optarg ? optarg : "";
        switch (c) {
        case 'q': o.mapq = atoi(a); break;
        case 'b': o.batch = atoi(a); break;
        case 't': o.thread = atoi(a); break;
        case 'i': o.input = a; break;
        case 'r': o.reference = a; break;
        case 'p': break;
        case 's': o.region = a; break;
        case 'g': o.group = a; break;
        case 'o': o.output = a; break;
        case 'a': o.maf = atof(a); break;
        case 8: o.rerun = true; break;
        case 7: o.load = true; break;
        case 6: o.keep_tmp = true; break;
        case 1: o.gpus = atoi(a); break;
        case 2: o.tile_sites = atoi(a); break;
        case 3: o.pack = true; break;
        case 4: o.no_packed = true; break;
        case 5: o.shard_samples = atoi(a); break;
        case 9: o.pending_mb = atoi(a); break;
        case 'v': o.verbose = true; break;
        default: die = true;
        }
    }
    if (die || o.input.empty() || o.output.empty()) { std::cerr << kUsage; return false; }
    return true;
}

// BaseVarUtils.h:59-75: "chr:S-E" -> (chr, S, E-1)
bool split_region(const std::string& rg, std::string& chr, long long& s, long long& e)
{
    const size_t c = rg.find(':');
    if (c == std::string::npos) return false;
    const std::string rest = rg.substr(c + 1);
    const size_t d = rest.find('-');
    if (d == std::string::npos) return false;
    chr = rg.substr(0, c);
    s = atoll(rest.substr(0, d).c_str());
    e = atoll(rest.substr(d + 1).c_str()) - 1;
    return true;
}

inline int base_code(char c)       // BASE_INT8_TABLE, src/BaseType.h:25-42
{
    switch (c) { case 'A': return 0; case 'C': return 1; case 'G': return 2; case 'T': return 3; default: return -1; }
}

std::string now_string()
{
    time_t t = time(nullptr);
    return std::string(ctime(&t));
}

inline long long now_ns()
{
    return std::chrono::duration_cast<std::chrono::nanoseconds>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

// ---- one tile in flight: the packed sites of a window, their records, and a completion flag ----------------
struct TileJob {
    explicit TileJob(int32_t n_samples) : pk(n_samples) {}
    Packer pk;
    // result of a tile: the SPLIT form (bvc_enqueue_split: a 72-byte core per site + the full records of the LRT-positive sites)
    // unless --group is on (its outputs hang off the full-record ticket)
    std::vector<bvc_site_core> core;
    std::vector<bvc_site_result> vcf;
    int32_t n_vcf = 0;
    std::vector<bvc_site_result> rec;      // full records (--group only)
    std::vector<int32_t> group_depth;      // [n_sites][n_groups][4]
    std::vector<double> group_af;
    std::mutex m;
    std::condition_variable cv;
    bool done = false;
    int rc = 0;
    std::string err;
};

// ---- one dispatcher thread per GPU: owns the context (not thread-safe), keeps two tickets in flight ----------
class GpuServer {
public:
    GpuServer(int device, const bvc_params& prm) : device_(device), prm_(prm) {}
    // the context is created by the dispatcher thread itself, so that CUDA start-up (a few hundred ms) overlaps the
    // window workers' first inflate / tokenise trips; a failure fails every job submitted to this GPU
    void start() { th_ = std::thread([this] { run(); }); }
    const std::string& init_error() const { return init_err_; }
    void submit(TileJob* j)
    {
        { std::lock_guard<std::mutex> lk(m_); q_.push_back(j); }
        cv_.notify_one();
    }
    void stop()
    {
        { std::lock_guard<std::mutex> lk(m_); stop_ = true; }
        cv_.notify_one();
        if (th_.joinable()) th_.join();
        if (ctx_) { launches_ = bvc_kernel_launch_count(ctx_); bvc_destroy(ctx_); ctx_ = nullptr; }
        free_stages();
    }
    long long launches() const { return launches_; }
    long long create_ns() const { return create_ns_; }
    long long busy_ns() const { return busy_ns_; }
    int tiles() const { return tiles_; }

private:
    struct Flight { TileJob* job; uint64_t ticket; int stage; };
    // Two page-locked staging sets, one per ticket in flight (allocated once, right after the context): a tile's arrays are
    // copied in (a few MB at memcpy speed), so bvc_enqueue_split's transfers run asynchronously from pinned memory, mapq / rpr
    // are read in place over PCIe at the LRT-positive sites only instead of being copied, and the results land in pinned
    // memory -- the data path bench.py's `e2e` measures.  (Page-locking every window's own arrays instead -- cudaHostRegister of
    // ~0.6 GB for 16 windows -- cost 0.65 s of a 1.4 s command.)  --group keeps the full-record path on the job's own arrays.
    struct Stage {
        uint8_t *pk8 = nullptr, *code = nullptr, *qual = nullptr, *mapq = nullptr, *rpr = nullptr, *ref = nullptr;
        uint32_t* off = nullptr;
        bvc_site_core* core = nullptr;
        bvc_site_result* vcf = nullptr;
    };
    bool alloc_stages()
    {
        if (prm_.n_groups > 0 || getenv("BVC_NO_HOST_STAGING")) return true;
        const size_t R = (size_t)prm_.max_entries + 64, S = (size_t)prm_.max_sites + 1;
        for (Stage& st : stage_) {
            st.pk8 = (uint8_t*)bvc_host_alloc(R); st.code = (uint8_t*)bvc_host_alloc(R); st.qual = (uint8_t*)bvc_host_alloc(R);
            st.mapq = (uint8_t*)bvc_host_alloc(R); st.rpr = (uint8_t*)bvc_host_alloc(R); st.ref = (uint8_t*)bvc_host_alloc(S);
            st.off = (uint32_t*)bvc_host_alloc((S + 1) * sizeof(uint32_t));
            st.core = (bvc_site_core*)bvc_host_alloc(S * sizeof(bvc_site_core));
            st.vcf = (bvc_site_result*)bvc_host_alloc(S * sizeof(bvc_site_result));
            if (!st.pk8 || !st.code || !st.qual || !st.mapq || !st.rpr || !st.ref || !st.off || !st.core || !st.vcf) return false;
        }
        staged_ = true;
        return true;
    }
    void free_stages()
    {
        for (Stage& st : stage_) {
            bvc_host_free(st.pk8); bvc_host_free(st.code); bvc_host_free(st.qual); bvc_host_free(st.mapq); bvc_host_free(st.rpr);
            bvc_host_free(st.ref); bvc_host_free(st.off); bvc_host_free(st.core); bvc_host_free(st.vcf);
            st = Stage();
        }
    }
    void finish(const Flight& f)
    {
        const long long tb = now_ns();
        int rc = prm_.n_groups > 0 ? bvc_wait(ctx_, f.ticket) : bvc_wait_split(ctx_, f.ticket, &f.job->n_vcf);
        std::string err = rc == BVC_OK ? "" : bvc_last_error(ctx_);
        if (rc == BVC_OK && staged_ && f.stage >= 0) {               // results home from the staging set
            const Stage& st = stage_[f.stage];
            memcpy(f.job->core.data(), st.core, (size_t)f.job->pk.n_sites() * sizeof(bvc_site_core));
            if (f.job->n_vcf > 0) memcpy(f.job->vcf.data(), st.vcf, (size_t)f.job->n_vcf * sizeof(bvc_site_result));
        }
        if (rc == BVC_OK && prm_.n_groups > 0 && f.job->pk.n_sites() > 0) {
            rc = bvc_fetch_groups(ctx_, f.ticket, f.job->group_depth.data(), f.job->group_af.data());
            if (rc != BVC_OK) err = bvc_last_error(ctx_);
        }
        busy_ns_ += now_ns() - tb; ++tiles_;
        // notified under the lock: the window thread may free the job as soon as it sees `done`
        { std::lock_guard<std::mutex> lk(f.job->m); f.job->rc = rc; f.job->err = err; f.job->done = true; f.job->cv.notify_all(); }
    }
    void run()
    {
        const long long tc0 = now_ns();
        const int rc0 = bvc_create(&ctx_, device_, &prm_);
        int rc_stage = BVC_OK;
        if (rc0 == BVC_OK && !alloc_stages()) rc_stage = BVC_ERR_CUDA;
        create_ns_ = now_ns() - tc0;
        if (rc0 != BVC_OK || rc_stage != BVC_OK) {
            init_err_ = std::string("GPU ") + std::to_string(device_) + ": " + (rc0 != BVC_OK ? bvc_last_error(nullptr) : "cannot allocate the page-locked staging buffers");
            for (;;) {                                         // fail whatever is submitted until stop()
                TileJob* j = nullptr;
                {
                    std::unique_lock<std::mutex> lk(m_);
                    cv_.wait(lk, [this] { return stop_ || !q_.empty(); });
                    if (q_.empty()) return;
                    j = q_.front(); q_.pop_front();
                }
                { std::lock_guard<std::mutex> lk(j->m); j->rc = rc0 != BVC_OK ? rc0 : rc_stage; j->err = init_err_; j->done = true; j->cv.notify_all(); }
            }
        }
        std::deque<Flight> fly;
        for (;;) {
            TileJob* j = nullptr;
            {
                std::unique_lock<std::mutex> lk(m_);
                if (fly.empty()) cv_.wait(lk, [this] { return stop_ || !q_.empty(); });
                if (!q_.empty()) { j = q_.front(); q_.pop_front(); }
                else if (stop_ && fly.empty()) return;
            }
            // group outputs belong to the most recent ticket of a slot: with groups, retire before re-using a slot
            const size_t max_fly = prm_.n_groups > 0 ? 1 : 2;
            if (j) {
                while (fly.size() >= max_fly) { finish(fly.front()); fly.pop_front(); }
                bvc_tile t;
                j->pk.tile(&t);
                if (prm_.n_groups == 0) t.sample = nullptr;
                uint64_t tk = 0;
                const long long tb = now_ns();
                int stage = -1, rc;
                if (prm_.n_groups > 0) rc = bvc_enqueue(ctx_, &t, j->rec.data(), &tk);
                else if (!staged_) rc = bvc_enqueue_split(ctx_, &t, j->core.data(), j->vcf.data(), (int32_t)j->vcf.size(), &tk);
                else {
                    stage = (int)(n_enq_++ & 1);                     // at most two tickets in flight: the older one's set is free again
                    Stage& st = stage_[stage];
                    const size_t S = (size_t)t.n_sites, R = t.n_sites > 0 ? (size_t)t.entry_off[S] : 0;
                    memcpy(st.ref, t.ref_base, S); memcpy(st.off, t.entry_off, (S + 1) * sizeof(uint32_t));
                    memcpy(st.mapq, t.mapq, R); memcpy(st.rpr, t.rpr, R);
                    if (t.packed8) { memcpy(st.pk8, t.packed8, R); t.packed8 = st.pk8; t.code = nullptr; t.qual = nullptr; }
                    else { memcpy(st.code, t.code, R); memcpy(st.qual, t.qual, R); t.code = st.code; t.qual = st.qual; }
                    t.ref_base = st.ref; t.entry_off = st.off; t.mapq = st.mapq; t.rpr = st.rpr;
                    rc = bvc_enqueue_split(ctx_, &t, st.core, st.vcf, prm_.max_sites, &tk);
                }
                busy_ns_ += now_ns() - tb;
                if (rc != BVC_OK) {
                    { std::lock_guard<std::mutex> lk(j->m); j->rc = rc; j->err = bvc_last_error(ctx_); j->done = true; j->cv.notify_all(); }
                } else fly.push_back({j, tk, stage});
            } else if (!fly.empty()) { finish(fly.front()); fly.pop_front(); }
        }
    }
    int device_;
    bvc_params prm_;
    Stage stage_[2];
    bool staged_ = false;
    unsigned n_enq_ = 0;
    bvc_ctx* ctx_ = nullptr;
    std::thread th_;
    std::mutex m_;
    std::condition_variable cv_;
    std::deque<TileJob*> q_;
    bool stop_ = false;
    long long launches_ = 0;
    std::string init_err_;
    long long create_ns_ = 0, busy_ns_ = 0;          // context creation; time inside bvc_enqueue / bvc_wait / bvc_fetch_groups
    int tiles_ = 0;
};

struct Shared {
    Options opt;
    std::string chr, refseq;
    long long rg_s = 0;
    std::vector<int32_t> pv;
    int32_t N = 0, nb = 0, tile_sites = 0;
    bool packed_in = false;                 // the windows are read from packed tile files (PREFIX.tmp.thread.{t}/packed.bvc)
    std::string sams;                       // sample names, tab-separated (with --pack: stored in every packed file)
    int64_t tile_entries = 0;
    int max_jobs = 2;                       // tiles a window may hold (filled, in flight or being filled): 2 + what --pending_mb allows
    std::atomic<int> max_jobs_used{0};
    int n_groups = 0;
    std::vector<std::string> group_names;
    std::vector<const char*> group_name_ptrs;
    std::string headvcf, headcvg;
    std::atomic<bool> failed{false};
    std::mutex err_m;
    std::string err;
    void fail(const std::string& e) { std::lock_guard<std::mutex> lk(err_m); if (!failed.exchange(true)) err = e; }
    // --verbose: thread-time per host phase, summed over the window workers (ns)
    std::atomic<long long> ns_read{0}, ns_pack{0}, ns_wait{0}, ns_format{0}, ns_write{0};
};



// ---- one thread per window: read + tokenise -> tile -> GPU -> format -> BGZF (in memory) ---------------------
void window_worker(Shared* sh, GpuServer* gpu, int ithread, std::string* vcf_out, std::string* cvg_out)
{
    const Options& o = sh->opt;
    const int T = o.thread;
    std::vector<std::unique_ptr<BgzfReader>> in;
    std::vector<std::string> names;
    const std::string dir = o.output + ".tmp.thread." + std::to_string(ithread);
    PackedIo pin, pout;                                      // packed tiles in (instead of the text files) / out (--pack)
    std::string line;
    if (sh->packed_in) {
        std::string e;
        if (!pin.open(dir + "/packed.bvc", &e)) { sh->fail(e); return; }
    } else {
        for (int j = 0; j < sh->nb; ++j) {
            names.push_back(dir + "/batch." + std::to_string(j));
            in.emplace_back(new BgzfReader());
            std::string e;
            if (!in.back()->open(names.back(), &e)) { sh->fail(e); return; }
        }
        for (auto& r : in) r->getline(line);                 // sample-name lines (header is built by the caller)
    }
    if (o.pack) {
        std::string e;
        if (!pout.create(dir + "/packed.bvc.tmp", sh->N, sh->sams, &e)) { sh->fail(e); return; }
    }
    BgzfWriter wv, wc;
    wv.open_memory(); wc.open_memory();
    if (ithread == 0) { wv.write(sh->headvcf); wc.write(sh->headcvg); }
    std::cerr << "begin to load data and run basetype" << std::endl;

    const size_t P = sh->pv.size();
    const size_t window = P % (size_t)T + P / (size_t)T;     // src/BaseVarC.cpp:399
    size_t it = std::min(P, (size_t)ithread * window);
    const size_t it_end = (ithread == T - 1) ? P : std::min(P, (size_t)(ithread + 1) * window);

    // Jobs of this window: `fly` = submitted and not yet formatted, in site order; `idle` = free to fill.  Two jobs are the steady
    // state (the GPU is far ahead of the packer); while the GPU is BEHIND -- CUDA start-up takes 0.6-1.0 s, a whole window's worth
    // of packing for a small job -- the worker does not wait for its oldest job but takes another one, up to sh->max_jobs (the
    // host-memory budget --pending_mb split over the windows), and keeps packing.
    std::vector<std::unique_ptr<TileJob>> all_jobs;
    std::deque<TileJob*> fly;
    std::vector<TileJob*> idle;
    auto new_job = [&]() -> TileJob* {
        all_jobs.emplace_back(new TileJob(sh->N));
        TileJob& j = *all_jobs.back();
        // text input: fixed capacity (a tile is flushed before it can outgrow it), no reallocation while a window is packed;
        // packed input: PackedIo::next_tile sizes the arrays to the tile it reads
        if (!sh->packed_in) j.pk.reserve((size_t)sh->tile_entries + 2 * (size_t)sh->N + 64);
        if (sh->n_groups == 0) { j.core.resize((size_t)sh->tile_sites + 1); j.vcf.resize((size_t)sh->tile_sites + 1); }
        return &j;
    };
    // the packers share ONE stale-token state in the reference (a single `ai` per window): hand it over by
    // packing strictly in site order and copying the state forward (Packer::adopt_state)
    int count = 0;
    long long t_read = 0, t_pack = 0, t_wait = 0, t_format = 0, t_write = 0;
    std::string vline, cline;
    // whatever way this function is left, no job may still be queued or in flight in the dispatcher: it writes the job's
    // records, status and condition variable, and bvc_enqueue's device-to-host copies target the job's arrays
    struct PendingGuard {
        std::deque<TileJob*>* fly;
        ~PendingGuard()
        {
            for (TileJob* j : *fly) { std::unique_lock<std::mutex> lk(j->m); j->cv.wait(lk, [&] { return j->done; }); }
            fly->clear();
        }
    } guard{&fly};

    auto drain = [&]() -> bool {                              // wait for the oldest job in flight, format its sites, release it
        TileJob& j = *fly.front();
        const long long tw = now_ns();
        { std::unique_lock<std::mutex> lk(j.m); j.cv.wait(lk, [&] { return j.done; }); }
        t_wait += now_ns() - tw;
        fly.pop_front();
        idle.push_back(&j);
        if (j.rc != BVC_OK) { sh->fail("GPU call failed: " + j.err); return false; }
        bvc_tile t;
        j.pk.tile(&t);
        const int32_t* pos = j.pk.positions();
        const int G = sh->n_groups;
        bvc_site_result full;
        for (int32_t s = 0; s < j.pk.n_sites(); ++s) {
            const bvc_site_result* r;
            if (G) r = &j.rec[(size_t)s];
            else {
                // split result: the full record of an LRT-positive site is in vcf[], every other site's VCF half is zero
                const bvc_site_core& c = j.core[(size_t)s];
                if (c.lrt_ok) {
                    if (c.vcf_index < 0 || c.vcf_index >= j.n_vcf) { sh->fail("GPU call returned an inconsistent split result"); return false; }
                    r = &j.vcf[(size_t)c.vcf_index];
                } else {
                    memset(&full, 0, sizeof full);
                    full.cvg_fs = c.cvg_fs; full.cvg_sor = c.cvg_sor; full.n_entries = c.n_entries;
                    for (int b = 0; b < 4; ++b) { full.depth[b] = c.depth[b]; full.cvg_strand[b] = c.cvg_strand[b]; }
                    full.em_fits = c.em_fits; full.em_passes = c.em_passes; full.flags = c.flags;
                    full.cvg_alt_base = c.cvg_alt_base; full.n_alt = c.n_alt;
                    full.alt_bases[0] = full.alt_bases[1] = full.alt_bases[2] = -1;
                    r = &full;
                }
            }
            const int32_t* gd = G ? &j.group_depth[(size_t)s * G * 4] : nullptr;
            const double* ga = G ? &j.group_af[(size_t)s * G * 4] : nullptr;
            const long long t0 = now_ns();
            if (r->lrt_ok) j.pk.expand_site(s);
            if (r->lrt_ok) format_vcf(vline, sh->chr.c_str(), pos[s], &t, s, r, sh->N, G, sh->group_name_ptrs.data(), gd, ga);
            format_cvg(cline, sh->chr.c_str(), pos[s], &t, s, r, &j.pk, G, gd);
            const long long t1 = now_ns();
            if (r->lrt_ok) wv.write(vline);
            wc.write(cline);
            t_format += t1 - t0; t_write += now_ns() - t1;
            if (!(++count % 1000)) std::cerr << "basetype completed " << count << " sites -- thread" << ithread << std::endl;
        }
        return true;
    };
    auto flush = [&](TileJob& j) {                            // hand a job to the GPU
        if (sh->n_groups) {
            j.rec.resize((size_t)j.pk.n_sites());
            j.group_depth.assign((size_t)j.pk.n_sites() * sh->n_groups * 4, 0);
            j.group_af.assign((size_t)j.pk.n_sites() * sh->n_groups * 4, 0.0);
        }
        j.done = false;
        fly.push_back(&j);
        gpu->submit(&j);
    };
    auto is_done = [](TileJob& j) { std::lock_guard<std::mutex> lk(j.m); return j.done; };
    // a job to fill next: an idle one; else the oldest in flight if the GPU has finished it (or no further job may be taken)
    auto take_job = [&]() -> TileJob* {
        while (idle.empty()) {
            if (!fly.empty() && (is_done(*fly.front()) || (int)all_jobs.size() >= std::max(2, sh->max_jobs))) { if (!drain()) return nullptr; }
            else return new_job();
        }
        TileJob* j = idle.back();
        idle.pop_back();
        return j;
    };

    // one tile into job k: sites of the window until the tile is full or the window ends.  1: more input may follow, 0: the
    // window is exhausted, -1: error (reported)
    auto fill_text = [&](TileJob& j) -> int {
        for (; it < it_end; ++it) {
            const int32_t p = sh->pv[it];
            const int rb = base_code(sh->refseq[(size_t)(p - sh->rg_s)]);
            j.pk.begin_site(p, (uint8_t)rb);
            for (auto& r : in) {
                const char* lp; size_t ln;
                const long long t0 = now_ns();
                const bool got = r->getline_view(&lp, &ln);
                const long long t1 = now_ns();
                t_read += t1 - t0;
                if (got) {
                    if (j.pk.add_line(lp, ln) != BVC_OK) { sh->fail(j.pk.error()); return -1; }
                    t_pack += now_ns() - t1;
                } else if (r->failed()) { sh->fail("corrupt batch file: " + r->error()); return -1; }
            }
            if (j.pk.end_site() == 0) j.pk.drop_last_site();  // no entries: the reference prints nothing (:442)
            if (j.pk.n_sites() >= sh->tile_sites || j.pk.n_entries() + sh->N > sh->tile_entries) { ++it; return 1; }
        }
        return 0;
    };
    auto fill_packed = [&](TileJob& j) -> int {
        std::string e;
        const long long t0 = now_ns();
        const int rc = pin.next_tile(j.pk, &e, sh->n_groups == 0);      // lazily: only the VCF columns of an LRT-positive site read code / qual / sample
        t_read += now_ns() - t0;
        if (rc < 0) sh->fail(e);
        return rc;
    };
    if (o.pack) {
        // --pack: the same read + tokenise loop, the tiles go to PREFIX.tmp.thread.{t}/packed.bvc instead of the GPU
        TileJob& j = *new_job();
        for (int more = 1; more > 0 && !sh->failed;) {
            more = fill_text(j);
            if (more < 0) return;
            std::string e;
            if (j.pk.n_sites() > 0 && !pout.add_tile(j.pk, &e)) { sh->fail(e); return; }
            j.pk.clear();                                      // keeps the stale-token state (Q1)
        }
        std::string e;
        if (!pout.finish(&e) || rename((dir + "/packed.bvc.tmp").c_str(), (dir + "/packed.bvc").c_str()) != 0) { sh->fail(e.empty() ? "cannot rename packed.bvc.tmp" : e); return; }
        sh->ns_read += t_read; sh->ns_pack += t_pack;
        return;
    }
    TileJob* prev = nullptr;
    for (int more = 1; more > 0 && !sh->failed;) {
        TileJob* j = take_job();
        if (!j) return;
        j->pk.clear();
        if (prev) j->pk.adopt_state(prev->pk);
        do {                                                   // (a tile whose sites are all empty is not submitted)
            more = sh->packed_in ? fill_packed(*j) : fill_text(*j);
            if (more < 0) return;
        } while (more > 0 && j->pk.n_sites() == 0 && !sh->failed);
        if (j->pk.n_sites() == 0) { idle.push_back(j); break; }
        flush(*j);
        prev = j;
    }
    if (sh->failed) return;                                   // (the guard waits for what is still in flight)
    while (!fly.empty()) if (!drain()) return;
    sh->max_jobs_used.store(std::max(sh->max_jobs_used.load(), (int)all_jobs.size()));
    { const long long t0 = now_ns(); wv.close(false); wc.close(false); t_write += now_ns() - t0; }
    sh->ns_read += t_read; sh->ns_pack += t_pack; sh->ns_wait += t_wait; sh->ns_format += t_format; sh->ns_write += t_write;
    *vcf_out = std::move(wv.memory());
    *cvg_out = std::move(wc.memory());
    if (!o.keep_tmp) {                                         // src/BaseVarC.cpp:458-462
        for (auto& n : names) remove(n.c_str());
        if (sh->packed_in) remove((dir + "/packed.bvc").c_str());
    }
}

#include "shard_rerun.inc"

}  // namespace
}  // namespace bvc_host

// =====================================================================================================
extern "C" int bvc_basetype_main(int argc, char** argv)
{
    using namespace bvc_host;
    Shared sh;
    Options& o = sh.opt;
    if (!parse_options(argc, argv, o)) return EXIT_FAILURE;
    std::string chr; long long rg_s = 0, rg_e = 0;
    if (!split_region(o.region, chr, rg_s, rg_e)) {
        std::cerr << "region must be feed with samtools-like format, i.e. chr:start-end" << std::endl;
        return EXIT_FAILURE;
    }
    if (o.reference.empty()) { std::cerr << "reference must be feed" << std::endl; return EXIT_FAILURE; }
    const clock_t ctb = clock();
    std::cout << "basetype start -- " << now_string();
    // CUDA start-up runs beside the file checks below (driver initialisation alone is ~0.1-0.2 s)
    const long long t_init0 = now_ns();
    auto devices = std::async(std::launch::async, [] {
        const int n = bvc_device_count();
        return std::make_pair(n, std::string(n > 0 ? "" : bvc_last_error(nullptr)));
    });
    // N = number of lines of the BAM list (src/BaseVarC.cpp:186-189); the BAMs themselves are never opened here
    {
        std::ifstream ib(o.input);
        if (!ib.is_open()) { std::cerr << "cannot open " << o.input << std::endl; return EXIT_FAILURE; }
        std::string l; int32_t n = 0;
        while (std::getline(ib, l)) ++n;
        sh.N = n;
    }
    if (sh.N <= 0 || o.thread <= 0 || o.batch <= 0) { std::cerr << "empty input list or bad --thread/--batch" << std::endl; return EXIT_FAILURE; }
    // reference bases of [S, E-1+1000) (src/BaseVarC.cpp:193-196 with both splitrg decrements) and the site list
    Fasta fa;
    std::string err;
    if (!fa.open(o.reference, &err) || !fa.fetch(chr, rg_s - 1, rg_e + 998, sh.refseq, &err)) { std::cerr << "ERROR: " << err << std::endl; return EXIT_FAILURE; }
    const long long buf = 1000;
    if ((long long)sh.refseq.size() <= buf) { std::cerr << "ERROR: region is empty or past the end of " << chr << std::endl; return EXIT_FAILURE; }
    for (size_t i = 0; i < sh.refseq.size() - (size_t)buf; ++i)
        if (base_code(sh.refseq[i]) >= 0) sh.pv.push_back((int32_t)(i + rg_s));
    sh.chr = chr; sh.rg_s = rg_s;
    const int T = o.thread;
    sh.nb = 1 + (sh.N - 1) / o.batch;
    if (T > 1 && (size_t)(T - 1) * (sh.pv.size() % T + sh.pv.size() / T) > sh.pv.size()) {
        std::cerr << "ERROR: --thread " << T << " does not partition " << sh.pv.size() << " sites (the reference reads past its site list here)" << std::endl;
        return EXIT_FAILURE;
    }
    // packed tile files (written by --pack or by a --load patched as INTEGRATION.md shows): when every window
    // has a complete one for this cohort they are read instead of the text batch files -- no inflate, no tokeniser
    if (!o.no_packed && !o.pack) {
        int npk = 0;
        for (int i = 0; i < T; ++i) if (PackedIo::is_complete(o.output + ".tmp.thread." + std::to_string(i) + "/packed.bvc", sh.N)) ++npk;
        sh.packed_in = npk == T;
    }
    // batch files: all T x nb must be complete BGZF files (src/BaseVarC.cpp:214-233)
    int ngz = 0;
    for (int j = 0; j < sh.nb && !sh.packed_in; ++j)
        for (int i = 0; i < T; ++i)
            if (bgzf_is_complete(o.output + ".tmp.thread." + std::to_string(i) + "/batch." + std::to_string(j))) ++ngz;
    if (!sh.packed_in && ngz != T * sh.nb) {
        std::cerr << "ERROR: " << ngz << " of " << T * sh.nb << " batch files are complete under " << o.output
                  << ".tmp.thread.*; BAM decoding stays with the reference: run `BaseVarC basetype --load` with the same -t/-b first" << std::endl;
        return EXIT_FAILURE;
    }
    std::cout << "basetype loading done -- " << now_string();
    if (o.load) return EXIT_SUCCESS;
    // with fewer window threads (or ranks of --shard_samples) than cores, the spare cores inflate the BGZF
    // blocks of the text batch files ahead of the tokenisers and deflate the windows' output blocks (one core is left to the GPU
    // dispatcher)
    struct PoolGuard { ~PoolGuard() { InflatePool::instance().stop(); } } pool_guard;
    {
        cpu_set_t cs;
        const int ncpu = sched_getaffinity(0, sizeof cs, &cs) == 0 ? CPU_COUNT(&cs) : (int)std::thread::hardware_concurrency();
        const int readers = o.shard_samples > 0 ? o.shard_samples : T;
        int n = ncpu - readers - (o.pack ? 0 : 1);
        if (const char* e = getenv("BVC_INFLATE_THREADS")) n = atoi(e);
        InflatePool::instance().start(n);
        if (o.verbose) std::cout << "basetype inflate pool : " << InflatePool::instance().threads() << " threads beside " << readers << " reading threads on " << ncpu << " cores" << std::endl;
    }

    // sample names = first line of every batch file of window 0, concatenated (src/BaseVarC.cpp:333-338)
    std::string sams;
    int64_t packed_sites = 0, packed_entries = 0;
    if (sh.packed_in) {
        for (int i = 0; i < T; ++i) {
            PackedIo r;
            if (!r.open(o.output + ".tmp.thread." + std::to_string(i) + "/packed.bvc", &err)) { std::cerr << "ERROR: " << err << std::endl; return EXIT_FAILURE; }
            if (i == 0) sams = r.names();
            packed_sites = std::max<int64_t>(packed_sites, r.header().max_sites);
            packed_entries = std::max<int64_t>(packed_entries, (int64_t)r.header().max_entries);
        }
        std::cerr << "reading packed tile files (" << o.output << ".tmp.thread.*/packed.bvc) instead of the text batch files" << std::endl;
    } else {
        for (int j = 0; j < sh.nb; ++j) {
            BgzfReader r; std::string l;
            if (!r.open(o.output + ".tmp.thread.0/batch." + std::to_string(j), &err)) { std::cerr << "ERROR: " << err << std::endl; return EXIT_FAILURE; }
            if (r.getline(l)) sams += l;
        }
        if (!sams.empty()) sams.pop_back();
    }
    sh.sams = sams;
    // population groups (src/BaseVarC.cpp:340-370): std::map order = sorted group names
    std::vector<int32_t> sample_group((size_t)sh.N, -1);
    sh.headcvg = kCvgHeader; sh.headvcf = kVcfHeader;
    if (!o.group.empty()) {
        std::ifstream ifg(o.group);
        std::unordered_map<std::string, std::string> popg;
        std::string id, grp;
        while (ifg >> id >> grp) popg.insert({id, grp});
        std::map<std::string, std::vector<int>> idx;
        std::istringstream iss(sams);
        int i = -1;
        while (std::getline(iss, id, '\t')) { ++i; auto f = popg.find(id); if (f != popg.end()) idx[f->second].push_back(i); }
        for (auto& kv : idx) {
            for (int s : kv.second) if (s < sh.N) sample_group[(size_t)s] = (int32_t)sh.group_names.size();
            sh.group_names.push_back(kv.first);
            sh.headcvg += "\t" + kv.first;
            sh.headvcf += "##INFO=<ID=" + kv.first + "_AF,Number=A,Type=Float,Description=\"Allele frequency in the " + kv.first + " populations calculated based on LRT.[0,1]\">\n";
        }
        sh.n_groups = (int)sh.group_names.size();
        for (auto& g : sh.group_names) sh.group_name_ptrs.push_back(g.c_str());
    }
    for (auto& c : fa.contigs()) sh.headvcf += "##contig=<ID=" + c.name + ",length=" + std::to_string(c.len) + ">\n";
    sh.headvcf += "##reference=file://" + o.reference + "\n";
    sh.headvcf += "#CHROM\tPOS\tID\tREF\tALT\tQUAL\tFILTER\tINFO\tFORMAT\t" + sams + "\n";
    sh.headcvg += "\n";

    if (o.pack) {
        // text batch files -> packed tile files, one per window; no GPU involved
        sh.tile_sites = o.tile_sites > 0 ? o.tile_sites : (int32_t)std::max<int64_t>(64, std::min<int64_t>(2048, ((int64_t)8 << 20) / sh.N));
        sh.tile_entries = std::max<int64_t>((int64_t)sh.tile_sites * sh.N, 2 * (int64_t)sh.N);
        std::vector<std::string> none((size_t)T);
        std::vector<std::thread> packers;
        for (int i = 0; i < T; ++i) packers.emplace_back(window_worker, &sh, nullptr, i, &none[(size_t)i], &none[(size_t)i]);
        for (auto& w : packers) w.join();
        devices.wait();
        if (sh.failed) { std::cerr << "ERROR: " << sh.err << std::endl; return EXIT_FAILURE; }
        std::cout << "basetype packing done -- " << now_string();
        return EXIT_SUCCESS;
    }
    // tile geometry and one dispatcher per GPU
    const auto dev = devices.get();
    const int n_dev = dev.first;
    if (const char* e = getenv("BVC_GPUS")) if (o.gpus <= 0) o.gpus = atoi(e);
    if (n_dev <= 0) { std::cerr << "ERROR: no CUDA device: the basetype calling core has no CPU path (" << dev.second << ")" << std::endl; return EXIT_FAILURE; }
    if (o.shard_samples > 0) {
        if (sh.packed_in) { std::cerr << "ERROR: --shard_samples reads the text batch files (a packed tile file holds whole-cohort tiles): add --no_packed" << std::endl; return EXIT_FAILURE; }
        const int rc = run_sharded_windows(sh, o.shard_samples, n_dev);
        if (rc == EXIT_SUCCESS) {
            std::cout << "basetype computing done -- " << now_string();
            std::cout << "basetype elapsed cpu secs : " << double(clock() - ctb) / CLOCKS_PER_SEC << std::endl;
            std::cout << "basetype done" << std::endl;
        }
        return rc;
    }
    const int G = std::max(1, std::min(std::min(o.gpus > 0 ? o.gpus : n_dev, n_dev), T));
    // Tiles of ~2,048 sites (fewer for wide cohorts: <= 8 M entries): with one worker per window feeding it, the GPU is
    // never the bottleneck, so small tiles cost nothing there and let packing, kernels and formatting overlap inside a
    // window; the context's buffers (sized by max_entries) stay small, which keeps its creation short.
    sh.tile_sites = o.tile_sites > 0 ? o.tile_sites : (int32_t)std::max<int64_t>(64, std::min<int64_t>(2048, ((int64_t)8 << 20) / sh.N));
    sh.tile_entries = std::max<int64_t>((int64_t)sh.tile_sites * sh.N, 2 * (int64_t)sh.N);
    if (sh.packed_in) { sh.tile_sites = (int32_t)std::max<int64_t>(1, packed_sites); sh.tile_entries = std::max<int64_t>(packed_entries, 2 * (int64_t)sh.N); }
    {   // a job = the packer's six arrays (9 B per entry of capacity) + the split result arrays
        const double job_bytes = 9.0 * (double)(sh.tile_entries + 2 * (int64_t)sh.N + 64) + (double)sh.tile_sites * (sizeof(bvc_site_core) + sizeof(bvc_site_result));
        const double per_window = (double)std::max(0, o.pending_mb) * 1048576.0 / T;
        sh.max_jobs = (int)std::max(2.0, std::min(64.0, per_window / job_bytes));
    }
    bvc_params prm;
    bvc_default_params(&prm, sh.N, o.maf);
    prm.max_sites = sh.tile_sites;
    prm.max_entries = sh.tile_entries;
    prm.n_groups = sh.n_groups;
    prm.sample_group = sh.n_groups ? sample_group.data() : nullptr;
    std::vector<std::unique_ptr<GpuServer>> gpus;
    for (int g = 0; g < G; ++g) { gpus.emplace_back(new GpuServer(g, prm)); gpus.back()->start(); }
    const long long t_init1 = now_ns();
    // windows [g*T/G, (g+1)*T/G) -> GPU g
    std::vector<std::string> vcf((size_t)T), cvg((size_t)T);
    std::vector<std::thread> workers;
    for (int i = 0; i < T; ++i) workers.emplace_back(window_worker, &sh, gpus[(size_t)((long long)i * G / T)].get(), i, &vcf[(size_t)i], &cvg[(size_t)i]);
    for (auto& w : workers) w.join();
    const long long t_work1 = now_ns();
    long long launches = 0;
    long long create_ns = 0, busy_ns = 0; int n_tiles = 0;
    for (auto& s : gpus) { s->stop(); launches += s->launches(); create_ns = std::max(create_ns, s->create_ns()); busy_ns += s->busy_ns(); n_tiles += s->tiles(); }
    const long long t_stop1 = now_ns();
    for (auto& s : gpus) if (!s->init_error().empty()) { std::cerr << "ERROR: " << s->init_error() << std::endl; return EXIT_FAILURE; }
    if (sh.failed) { std::cerr << "ERROR: " << sh.err << std::endl; return EXIT_FAILURE; }
    // ordered concatenation of the windows' BGZF blocks + one EOF block (replaces the re-read merge of :268-295)
    for (int k = 0; k < 2; ++k) {
        const std::string path = o.output + (k ? ".cvg.gz" : ".vcf.gz");
        FILE* f = fopen(path.c_str(), "wb");
        if (!f) { std::cerr << "ERROR: cannot write " << path << std::endl; return EXIT_FAILURE; }
        for (int i = 0; i < T; ++i) { const std::string& s = k ? cvg[(size_t)i] : vcf[(size_t)i]; fwrite(s.data(), 1, s.size(), f); }
        fwrite(BgzfWriter::kEofBlock, 1, 28, f);
        if (fclose(f) != 0) std::cerr << "warning: file cannot be closed" << std::endl;
    }
    std::cout << "merge subfiles done" << std::endl;
    if (!o.keep_tmp) for (int i = 0; i < T; ++i) rmdir((o.output + ".tmp.thread." + std::to_string(i)).c_str());
    std::cout << "basetype computing done -- " << now_string();
    std::cout << "basetype elapsed cpu secs : " << double(clock() - ctb) / CLOCKS_PER_SEC << std::endl;
    if (o.verbose) {
        std::cout << "basetype gpu kernels launched : " << launches << " on " << G << " GPU(s)" << std::endl;
        char b[512];
        snprintf(b, sizeof b, "basetype host phases : checks beside CUDA start-up %.3f s, windows (context creation inside) %.3f s wall on %d threads, "
                 "context teardown %.3f s, output files %.3f s; context creation %.3f s, %d tiles, %.3f s inside enqueue / wait; thread-seconds: inflate+getline %.3f, "
                 "tokenise+pack %.3f, wait for gpu %.3f, format %.3f, bgzf deflate %.3f; tiles held per window: up to %d of %d allowed",
                 (t_init1 - t_init0) * 1e-9, (t_work1 - t_init1) * 1e-9, T, (t_stop1 - t_work1) * 1e-9, (now_ns() - t_stop1) * 1e-9, create_ns * 1e-9, n_tiles, busy_ns * 1e-9, sh.ns_read * 1e-9, sh.ns_pack * 1e-9, sh.ns_wait * 1e-9,
                 sh.ns_format * 1e-9, sh.ns_write * 1e-9, sh.max_jobs_used.load(), sh.max_jobs);
        std::cout << b << std::endl;
    }
    std::cout << "basetype done" << std::endl;
    return EXIT_SUCCESS;
}
