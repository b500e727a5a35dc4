// Command-line front end with the reference's flags, defaults, positional arguments, exit codes
// and stdout narrative (reference src/Main.cpp), driving the GPU-backed classes.  The reference's
// own src/Main.cpp also compiles unchanged against telomeri_b200/host/include;
// this file exists so the binary can be built where the reference tree is absent.
#include <cstring>
#include <iostream>
#include <map>

#include <PathManager.hpp>
#include <Scaffolder.hpp>
#include <Stopwatch.hpp>

namespace {

using Pair = std::pair<const OverlapGraph::Node *, const OverlapGraph::Node *>;

[[noreturn]] void usage_and_exit() {
    std::cerr << "Please provide a path to read-read and contig-read files!\n"
              << "Usage: hera [option...] <read-read_overlap_file>.paf <contig-read_overlap_file>.paf "
              << "<reads_file>.fasta <contigs_file>.fasta <output_file>.fasta\n\n"
              << "Options:\n"
              << "  --filter-avg | --filter-min | --filter-max | --filter-sum   length mode of the overlap filter (default avg)\n"
              << "  --min-oll N   minimum overlap length (default 0)\n"
              << "  --min-olp F   minimum overlap length / total length (default 0.0)\n"
              << "  --max-ohl N   maximum overhang length (default unlimited)\n"
              << "  --max-ohp F   maximum overhang length / overlap length (default 1.0)\n"
              << "  --rb-att N    Monte-Carlo rebuild attempts per anchor (default 2000)\n"
              << "  --bt-att N    backtrack attempts per walk (default 60)\n"
              << "  --len-thr N   path length threshold (default 4700000)\n"
              << "  --w-size N    window size in path length (default 1000)\n"
              << "  --r-thr F     valley/peak ratio threshold (default 0.9)\n\n"
              << "Percentages are in range [0.0, 1.0].\n" << std::endl;
    std::exit(1);
}

ulong parse_count(const char *s) {
    try {
        const long v = std::stol(s);
        if (v < 0) { std::cerr << "Value must be positive: " << s << std::endl; std::exit(1); }
        return static_cast<ulong>(v);
    } catch (std::invalid_argument &) {
        std::cerr << "Invalid int value: " << s << std::endl;
    } catch (std::out_of_range &) {
        std::cerr << "Provided value is outside of int range: " << s << std::endl;
    }
    std::exit(1);
}

float parse_fraction(const char *s) {
    try {
        const float v = std::stof(s);
        if (v < 0.0f || v > 1.0f) { std::cerr << "Value must be in range [0.0, 1.0]: " << s << std::endl; std::exit(1); }
        return v;
    } catch (std::invalid_argument &) {
        std::cerr << "Invalid float value: " << s << std::endl;
    } catch (std::out_of_range &) {
        std::cerr << "Provided value is outside of float range: " << s << std::endl;
    }
    std::exit(1);
}

}  // namespace

int main(int argc, char **argv) {
    OverlapGraph::FilterParameters fp = {OverlapGraph::AVG, 0, 0.0f, ULONG_MAX, 1.0f};
    PathManager::Parameters pmp = {2000, 60, 4700000l, 1000ul, 0.9f};
    const char *mode_string = "AVG";
    if (argc < 6) usage_and_exit();
    for (int i = 1; i < argc - 5; i++) {
        const std::string a = argv[i];
        auto value = [&]() -> const char * { return i + 1 < argc - 5 ? argv[++i] : nullptr; };
        if (a == "--filter-avg") { fp.mode = OverlapGraph::AVG; mode_string = "AVG"; }
        else if (a == "--filter-min") { fp.mode = OverlapGraph::MIN; mode_string = "MIN"; }
        else if (a == "--filter-max") { fp.mode = OverlapGraph::MAX; mode_string = "MAX"; }
        else if (a == "--filter-sum") { fp.mode = OverlapGraph::SUM; mode_string = "SUM"; }
        else if (a == "--min-oll" || a == "--min-olp" || a == "--max-ohl" || a == "--max-ohp" || a == "--rb-att" || a == "--bt-att" ||
                 a == "--len-thr" || a == "--w-size" || a == "--r-thr") {
            const char *v = value();
            if (!v) break;  // option without a value before the positionals: the reference leaves it pending
            if (a == "--min-oll") fp.min_overlap_length = parse_count(v);
            else if (a == "--min-olp") fp.min_overlap_percentage = parse_fraction(v);
            else if (a == "--max-ohl") fp.max_overhang_length = parse_count(v);
            else if (a == "--max-ohp") fp.max_overhang_percentage = parse_fraction(v);
            else if (a == "--rb-att") pmp.rebuild_attempts = static_cast<int>(parse_count(v));
            else if (a == "--bt-att") pmp.backtrack_attempts = static_cast<int>(parse_count(v));
            else if (a == "--len-thr") pmp.len_threshold = static_cast<long>(parse_count(v));
            else if (a == "--w-size") pmp.window_size = parse_count(v);
            else pmp.ratio_threshold = parse_fraction(v);
        } else {
            std::cerr << "Unknown option: " << a << std::endl;
            return 1;
        }
    }
    char *rr_file = argv[argc - 5], *cr_file = argv[argc - 4], *reads_file = argv[argc - 3], *contigs_file = argv[argc - 2], *output_file = argv[argc - 1];

    std::cout << "Filter:\n    Mode: " << mode_string << "\n    Min overlap length: " << fp.min_overlap_length
              << "\n    Min overlap percentage: " << fp.min_overlap_percentage << "\n    Max overhang length: " << fp.max_overhang_length
              << "\n    Max overhang percentage: " << fp.max_overhang_percentage << std::endl;
    std::cout << "Path construction:\n    Rebuild attempts: " << pmp.rebuild_attempts << "\n    Backtrack attempts: " << pmp.backtrack_attempts
              << "\n    Length threshold: " << pmp.len_threshold << "\n    Window size: " << pmp.window_size
              << "\n    Ratio threshold: " << pmp.ratio_threshold << "\n" << std::endl;

    Stopwatch timer;
    timer.start();
    OverlapGraph graph;
    graph.filter_params_ = fp;
    std::cout << "\nLoading overlaps..." << std::endl;
    if (!graph.load(rr_file, false) || !graph.load(cr_file, true)) return 1;
    std::cout << "Done (" << timer.lap() << "s)" << std::endl << graph.stats() << std::endl;

    std::cout << "Calculating paths..." << std::endl;
    PathManager pm;
    pm.params_ = pmp;
    pm.buildMonteCarlo(graph, Utils::Metrics::EXTENSION_SCORE);
    pm.buildMonteCarlo(graph, Utils::Metrics::OVERLAP_SCORE);
    pm.buildDeterministic(graph, Utils::Metrics::EXTENSION_SCORE);
    pm.buildDeterministic(graph, Utils::Metrics::OVERLAP_SCORE);
    pm.filterUnique();
    std::cout << "Done (" << timer.lap() << "s)" << std::endl << pm.stats() << std::endl;

    std::map<Pair, std::vector<const Path *>> between = pm.getPathsBetweenAnchors();
    std::map<Pair, std::vector<PathGroup>> groups;
    for (auto &kv : between) {
        std::cout << "====> Constructing groups for paths between anchor '" << kv.first.first->name << "' and anchor '" << kv.first.second->name
                  << "'..." << std::endl;
        groups[kv.first] = PathManager::constructGroups(kv.second, pm.params_);
        std::cout << "<==== Finished constructing groups for paths between anchor '" << kv.first.first->name << "' and anchor '"
                  << kv.first.second->name << "'!\n" << std::endl;
    }

    std::map<Pair, const Path *> consensus;
    long consensus_num = 0;
    for (auto &kv : groups) {
        const std::string &n1 = kv.first.first->name, &n2 = kv.first.second->name;
        std::cout << "====> Finding consensus path in each group between anchor '" << n1 << "' and anchor '" << n2 << "'..." << std::endl;
        std::vector<PathGroup *> with;
        for (PathGroup &pg : kv.second) {
            pg.discardNotFrequent();
            pg.calculateConsensusPath();
            pg.calculateValidPathNumber();
            if (pg.consensus) {
                std::cout << "Consensus length: " << pg.consensus->length() << "    Valid path number: " << pg.valid_path_number << std::endl;
                with.push_back(&pg);
            } else {
                std::cout << "No consensus sequence for group!\n" << std::endl;
            }
        }
        std::cout << "Done finding consensus path in each group.\n";
        const Path *chosen = nullptr;
        if (with.size() == 1) {
            chosen = with[0]->consensus;
        } else if (with.size() >= 2) {  // longest consensus first; from three groups on, the valid-path tournament (src/Main.cpp:306-332)
            std::sort(with.begin(), with.end(), [](const PathGroup *a, const PathGroup *b) { return a->consensus->length() > b->consensus->length(); });
            const PathGroup *keep = with[0];
            if (with.size() > 2)
                for (size_t i = 1; i < with.size(); i++)
                    if (2 * keep->valid_path_number <= with[i]->valid_path_number) keep = with[i];
            chosen = keep->consensus;
        }
        consensus[kv.first] = chosen;
        if (chosen && ++consensus_num) std::cout << "Final consensus (consensus among groups) length: " << chosen->length() << '\n';
        else std::cout << "Final consensus not found!\n";
        std::cout << "====> Done finding consensus between anchor '" << n1 << "' and anchor '" << n2 << "'!" << std::endl;
    }
    if (consensus_num == 0) {
        std::cerr << "No consensus found!" << std::endl;
        return 1;
    }

    std::cout << "Building the scaffold..." << std::endl;
    Path scaffold = pm.constructConsensusPath(between, consensus, 10);
    std::cout << "Done (" << timer.lap() << "s)" << std::endl;
    std::cout << "\nLoading files for scaffolding..." << std::endl;
    Scaffolder scaff(scaffold);
    if (!scaff.load(reads_file) || !scaff.load(contigs_file)) return 1;
    std::cout << "Done (" << timer.lap() << "s)" << std::endl;
    std::cout << "Writing result to: " << output_file << std::endl;
    if (!scaff.write(output_file)) return 1;
    std::cout << "Done (" << timer.lap() << "s)" << std::endl;
    std::cout << "Total time: " << timer.stop() << "s" << std::endl;
    return 0;
}
