// vga_map — harness CLI standing in for the Rust front end (no cargo in this image).  It mirrors
// the reference's two subcommands and their flags (src/subcommands/cli.yml, index_main.rs,
// map_main.rs) and drives the hot path through the C ABI exactly like the Rust shim would:
//
//   vga_map index -i graph.gfa -k 11 -o graph            (writes graph.vgidx; `.idx` needs Rust/bincode)
//   vga_map map   -i graph.vgidx -f reads.fa -o out [--also-align] [-n best] [-c min-anchors]
//                 [-g max-gap] [--gpus N] (writes out-chains.gaf and out-alignments.gaf, src/map.rs:135-184)
// --gpus N: one host thread per GPU; the index image is replicated by vga_index_broadcast (ONE ncclBroadcast), every thread maps
// a contiguous range of the reads, the GAF texts are concatenated in rank order (output order = input order).
//
// The reference's -G/--input-graph and -p/--poa-aligner are accepted and ignored: the graph is only
// needed there for the per-read subgraph export (src/map.rs:155-159), and the aligner is the CUDA POA.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

#include "../../include/vga.h"
#include "vga_host.h"

static int die(const std::string& m) { fprintf(stderr, "vga_map: %s\n", m.c_str()); return 1; }

static int write_file(const std::string& path, const char* data, uint64_t n) {
    FILE* f = fopen(path.c_str(), "wb");
    if (!f) return die("Couldn't create file " + path);
    fwrite(data, 1, n, f);
    fclose(f);
    return 0;
}

int main(int argc, char** argv) {
    if (argc < 2) return die("usage: vga_map index|map ...");
    std::string cmd = argv[1], in, reads_path, out, k = "11", max_gap = "1000", min_anchors = "3", best_n = "1";
    std::string max_furc = "100", max_deg = "100";
    bool also_align = false;
    int n_gpus = 1;
    for (int i = 2; i < argc; i++) {
        std::string a = argv[i];
        auto val = [&]() -> std::string { if (i + 1 >= argc) { die("missing value for " + a); exit(1); } return argv[++i]; };
        if (a == "-i" || a == "--input" || a == "--index") in = val();
        else if (a == "-f" || a == "--input-file") reads_path = val();
        else if (a == "-o" || a == "--out-prefix" || a == "--out") out = val();
        else if (a == "-k" || a == "--kmer-length") k = val();
        else if (a == "-e" || a == "--max-furcations") max_furc = val();
        else if (a == "-d" || a == "--max-degree") max_deg = val();
        else if (a == "-g" || a == "--max-gap-length") max_gap = val();
        else if (a == "-c" || a == "--chain-min-anchors") min_anchors = val();
        else if (a == "-n" || a == "--align-best-n") best_n = val();
        else if (a == "--also-align") also_align = true;
        else if (a == "--gpus") n_gpus = atoi(val().c_str());
        else if (a == "-G" || a == "--input-graph" || a == "-p" || a == "--poa-aligner" || a == "-t" || a == "--n-threads") val();
        else return die("unknown argument " + a);
    }
    if (cmd == "index") {
        if (in.empty() || out.empty()) return die("index needs -i <gfa> and -o <prefix>");
        vgah_index_t* ix = vgah_index_build_file(in.c_str(), (uint32_t)atoi(k.c_str()), strtoull(max_furc.c_str(), 0, 10),
                                                 strtoull(max_deg.c_str(), 0, 10));
        if (!ix) return die(vgah_last_error());
        std::string path = out.size() > 6 && out.substr(out.size() - 6) == ".vgidx" ? out : out + ".vgidx";
        if (vgah_index_save(ix, path.c_str()) != 0) return die(vgah_last_error());
        VgaIndexView v;
        vgah_index_view(ix, &v);
        fprintf(stderr, "Index with k=%u built correctly! Found %llu different kmers, which appear in %llu positions!\n", v.k,
                (unsigned long long)v.n_keys, (unsigned long long)(v.n_table - v.n_keys));
        vgah_index_free(ix);
        return 0;
    }
    if (cmd != "map") return die("unknown subcommand " + cmd);
    if (in.empty() || reads_path.empty()) return die("map needs -i <index> and -f <reads>");
    if (out.empty()) {      // same default as map_main.rs:20-28: input path without its extension
        size_t cut = (reads_path.size() >= 2 && reads_path.substr(reads_path.size() - 2) == "fa") ||
                             (reads_path.size() >= 5 && reads_path.substr(reads_path.size() - 5) == "fasta")
                         ? 3 : 4;
        out = reads_path.substr(0, reads_path.size() > cut ? reads_path.size() - cut : 0);
    }
    vgah_index_t* hix = vgah_index_load(in.c_str());
    if (!hix) return die(vgah_last_error());
    vgah_reads_t* reads = vgah_reads_from_file(reads_path.c_str());
    if (!reads) return die(vgah_last_error());
    VgaIndexView view;
    vgah_index_view(hix, &view);
    VgaParams P;
    vga_default_params(&P);
    P.max_gap = (uint32_t)atoi(max_gap.c_str()); P.min_anchors = (uint32_t)atoi(min_anchors.c_str());
    P.best_n = (uint32_t)atoi(best_n.c_str());
    if (!also_align) P.flags |= VGA_FLAG_NO_ALIGN;
    if (n_gpus < 1 || n_gpus > vga_device_count()) return die("--gpus outside 1.." + std::to_string(vga_device_count()));
    const uint32_t n_reads = vgah_reads_count(reads);
    const uint64_t* offs = vgah_reads_offsets(reads);
    const char* names = vgah_reads_names(reads);
    // rank r maps reads [r*n/N, (r+1)*n/N); name_at[r] = start of its first name in the '\n' separated list
    std::vector<const char*> name_at(n_gpus + 1, names);
    {
        const char* p = names;
        uint32_t line = 0;
        for (int r = 1; r <= n_gpus; r++) {
            const uint32_t lo = (uint32_t)((uint64_t)n_reads * r / n_gpus);
            while (line < lo && *p) { if (*p == '\n') line++; p++; }
            name_at[r] = p;
        }
    }
    unsigned char comm_id[VGA_COMM_ID_BYTES];
    std::vector<vga_index_t*> gix(n_gpus, nullptr);
    if (vga_init(0) != VGA_OK || vga_index_upload(&view, &gix[0]) != VGA_OK) return die(vga_last_error());
    if (n_gpus > 1 && vga_comm_unique_id(comm_id) != VGA_OK) return die(vga_last_error());
    std::vector<std::string> text0(n_gpus), text1(n_gpus), errs(n_gpus);
    std::vector<VgaStats> stats(n_gpus);
    auto worker = [&](int r) {
        auto fail = [&](const char* m) { errs[r] = m ? m : "error"; };
        if (vga_init(r) != VGA_OK) return fail(vga_last_error());
        if (vga_index_broadcast(&gix[r], comm_id, r, n_gpus) != VGA_OK) return fail(vga_last_error());
        const uint32_t lo = (uint32_t)((uint64_t)n_reads * r / n_gpus), hi = (uint32_t)((uint64_t)n_reads * (r + 1) / n_gpus);
        vga_results_t* res = nullptr;
        if (vga_map_batch(gix[r], vgah_reads_seqs(reads), offs + lo, hi - lo, &P, &res) != VGA_OK) return fail(vga_last_error());
        VgaResultsView rv;
        vga_results_view(res, &rv);
        const std::string my_names(name_at[r], name_at[r + 1]);
        for (int which = 0; which <= (also_align ? 1 : 0); which++) {
            uint64_t n = 0, failed = 0;
            char* text = vgah_gaf_raw(hix, &rv, my_names.c_str(), lo, vgah_reads_seqs(reads), offs + lo, which, 0, &n, &failed);
            if (!text) return fail(vgah_last_error());
            (which ? text1 : text0)[r].assign(text, n);
            vgah_free_str(text);
            if (failed) fprintf(stderr, "vga_map: %llu alignments of rank %d did not complete on the device (written as placeholders)\n",
                                (unsigned long long)failed, r);
        }
        vga_results_stats(res, &stats[r]);
        vga_results_free(res);
        vga_index_free(gix[r]);
    };
    {
        std::vector<std::thread> th;
        for (int r = 1; r < n_gpus; r++) th.emplace_back(worker, r);
        worker(0);
        for (auto& t : th) t.join();
    }
    for (int r = 0; r < n_gpus; r++) if (!errs[r].empty()) return die("rank " + std::to_string(r) + ": " + errs[r]);
    const bool is_gaf = out.size() > 4 && out.substr(out.size() - 4) == ".gaf";
    for (int which = 0; which <= (also_align ? 1 : 0); which++) {
        std::string all;
        for (int r = 0; r < n_gpus; r++) all += (which ? text1 : text0)[r];
        if (write_file(is_gaf ? out : out + (which ? "-alignments.gaf" : "-chains.gaf"), all.data(), all.size())) return 1;
    }
    VgaStats st = stats[0];
    for (int r = 1; r < n_gpus; r++) {
        st.n_reads += stats[r].n_reads; st.n_chains += stats[r].n_chains; st.n_alignments += stats[r].n_alignments;
        if (stats[r].ms_total > st.ms_total) st.ms_total = stats[r].ms_total;
    }
    fprintf(stderr, "Found %llu reads, %llu chains, %llu alignments (%.1f ms on the GPU path, %d GPU%s)\n", (unsigned long long)st.n_reads,
            (unsigned long long)st.n_chains, (unsigned long long)st.n_alignments, st.ms_total, n_gpus, n_gpus > 1 ? "s" : "");
    vgah_reads_free(reads);
    vgah_index_free(hix);
    return 0;
}
