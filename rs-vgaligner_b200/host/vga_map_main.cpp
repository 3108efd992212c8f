// vga_map — harness CLI standing in for the Rust front end (no cargo in this image).  It mirrors
// the reference's two subcommands and their flags (src/subcommands/cli.yml, index_main.rs,
// map_main.rs) and drives the hot path through the C ABI exactly like the Rust shim would:
//
//   vga_map index -i graph.gfa -k 11 -o graph            (writes graph.vgidx; `.idx` needs Rust/bincode)
//   vga_map map   -i graph.vgidx -f reads.fa -o out [--also-align] [-n best] [-c min-anchors]
//                 [-g max-gap]          (writes out-chains.gaf and out-alignments.gaf, src/map.rs:135-184)
//
// The reference's -G/--input-graph and -p/--poa-aligner are accepted and ignored: the graph is only
// needed there for the per-read subgraph export (src/map.rs:155-159), and the aligner is the CUDA POA.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>

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
    vga_index_t* gix = nullptr;
    if (vga_index_upload(&view, &gix) != VGA_OK) return die(vga_last_error());
    VgaParams P;
    vga_default_params(&P);
    P.max_gap = (uint32_t)atoi(max_gap.c_str()); P.min_anchors = (uint32_t)atoi(min_anchors.c_str());
    P.best_n = (uint32_t)atoi(best_n.c_str());
    if (!also_align) P.flags |= VGA_FLAG_NO_ALIGN;
    vga_results_t* res = nullptr;
    if (vga_map_batch(gix, vgah_reads_seqs(reads), vgah_reads_offsets(reads), vgah_reads_count(reads), &P, &res) != VGA_OK)
        return die(vga_last_error());
    VgaResultsView rv;
    vga_results_view(res, &rv);
    const bool is_gaf = out.size() > 4 && out.substr(out.size() - 4) == ".gaf";
    uint64_t n = 0;
    char* text = vgah_gaf(hix, &rv, reads, 0, &n);
    if (write_file(is_gaf ? out : out + "-chains.gaf", text, n)) return 1;
    vgah_free_str(text);
    if (also_align) {
        text = vgah_gaf(hix, &rv, reads, 1, &n);
        if (write_file(is_gaf ? out : out + "-alignments.gaf", text, n)) return 1;
        vgah_free_str(text);
    }
    VgaStats st;
    vga_results_stats(res, &st);
    fprintf(stderr, "Found %llu reads, %llu chains, %llu alignments (%.1f ms on the GPU path)\n", (unsigned long long)st.n_reads,
            (unsigned long long)st.n_chains, (unsigned long long)st.n_alignments, st.ms_total);
    vga_results_free(res);
    vga_index_free(gix);
    vgah_reads_free(reads);
    vgah_index_free(hix);
    return 0;
}
