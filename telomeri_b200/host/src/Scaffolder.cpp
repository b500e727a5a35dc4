// Scaffolder (reference src/Scaffolder.cpp:12-147): loads reads/contigs and stitches the scaffold
// sequence.  Host code; the stitching rules (inclusive substring bounds, complement without
// reversal on relative_strand, trimming when the next overlap starts before the previous one ends,
// one global reverse at the end) are the reference's, restated.
#include <Scaffolder.hpp>
#include <OverlapGraph.hpp>

#include "hera_b200.h"

#include <algorithm>
#include <cstring>
#include <fstream>
#include <iostream>
#include <sstream>
#include <stdexcept>

unsigned long Scaffolder::min_path_num_ = 10;

bool Scaffolder::load(const char *filename) {
    files_.emplace_back(filename);
    if (OverlapGraph::current_device()) {   // parsed on the GPU by write(); only check that it can be read
        std::ifstream probe(filename);
        if (!probe) {
            std::cerr << "Cannot open file: " << filename << std::endl;
            return false;
        }
        return true;
    }
    return load_host(filename);
}

bool Scaffolder::load_host(const char *filename) {
    const size_t n = std::strlen(filename);
    const bool fastq = n > 0 && filename[n - 1] == 'q';
    std::ifstream in(filename);
    if (!in) {
        std::cerr << "Cannot open file: " << filename << std::endl;
        return false;
    }
    std::string line;
    for (bool header = true; std::getline(in, line); header = !header) {
        if (header) {
            std::istringstream first_word(line);
            first_word >> line;
            names_.emplace_back(line.empty() ? line : line.substr(1));  // drop '>' / '@'
        } else {
            sequences_.push_back(line);
            if (fastq) {  // '+' line and qualities
                std::getline(in, line);
                std::getline(in, line);
            }
        }
    }
    return true;
}

std::string &Scaffolder::getSequenceFrom(const std::string &name) {
    const size_t i = static_cast<size_t>(std::find(names_.begin(), names_.end(), name) - names_.begin());
    return sequences_.at(i);  // the reference indexes unchecked; a missing name is an error here
}

std::string Scaffolder::substring(std::string &str, ulong start, ulong end) { return str.substr(start, end - start + 1); }

std::string &Scaffolder::reverse(std::string &str) {
    std::reverse(str.begin(), str.end());
    return str;
}

std::string Scaffolder::reverse(std::string str) {
    std::reverse(str.begin(), str.end());
    return str;
}

void Scaffolder::swapBases(std::string &str) {
    for (char &c : str) c = c == 'A' ? 'T' : c == 'T' ? 'A' : c == 'G' ? 'C' : c == 'C' ? 'G' : c;
}

bool Scaffolder::write(const char *filename) {
    hera_ctx *ctx = OverlapGraph::current_device();
    if (ctx && files_.size() == 2 && p_.nodes_.size() >= 2) {
        hera_scaffold_view v{};
        const int rc = hera_scaffold(ctx, files_[0].c_str(), files_[1].c_str(), min_path_num_, &v);
        // the reference dies of std::out_of_range here (substr / vector::at): keep that observable behaviour
        if (rc == 44 || rc == 45) throw std::out_of_range(hera_last_error(ctx));
        bool same = rc == 0 && v.nodes == p_.nodes_.size();
        for (uint64_t i = 0; same && i < v.nodes; i++) same = v.node[i] == p_.nodes_[i]->index;
        if (rc != 0 || !same) {
            std::cerr << "hera_scaffold: " << (rc ? hera_last_error(ctx) : "the library's chain of consensus paths differs from the given path") << std::endl;
            return false;
        }
        std::ofstream out(filename, std::ios::binary);
        if (!out) {
            std::cerr << "Cannot write file: " << filename << std::endl;
            return false;
        }
        for (uint64_t i = 0; i < v.logged; i++) std::cout << (i ? "-" : "") << p_.nodes_[v.logged_index[i]]->name;
        std::cout << std::endl;
        out.write(v.fasta, static_cast<std::streamsize>(v.fasta_bytes));
        return true;
    }
    if (names_.empty()) for (const std::string &f : files_) if (!load_host(f.c_str())) return false;   // files were only probed by load()
    std::ofstream out(filename);
    if (!out) {
        std::cerr << "Cannot write file: " << filename << std::endl;
        return false;
    }
    out << ">Resulting_HERA_scaffold" << std::endl;
    const size_t nn = p_.nodes_.size(), ne = p_.edges_.size();
    if (nn == 0) {
        std::cerr << "Scaffold is empty!" << std::endl;
        return false;
    }
    if (nn == 1) {
        std::cout << p_.nodes_[0]->name << std::endl;
        out << reverse(std::string(getSequenceFrom(p_.nodes_[0]->name)));
        return true;
    }
    std::string seq, tmp;
    std::cout << p_.nodes_[0]->name;
    tmp = getSequenceFrom(p_.nodes_[0]->name);
    seq += reverse(substring(tmp, static_cast<ulong>(p_.edges_[0]->t_end) + 1, tmp.size()));
    for (size_t i = 1; i < ne; i++) {
        const ulong start = p_.edges_[i]->t_end;    // where the next overlap ends on this node
        const ulong end = p_.edges_[i - 1]->q_end;  // where the previous overlap ends on this node
        if (start < end) {
            tmp = getSequenceFrom(p_.nodes_[i]->name);
            if (p_.edges_[i - 1]->relative_strand) swapBases(tmp);
            std::cout << '-' << p_.nodes_[i]->name;
            seq += reverse(substring(tmp, start + 1, end));
        } else if (start > end) {
            seq = seq.substr(0, static_cast<ulong>(seq.length()) - (start - end));  // unsigned wrap keeps everything, like the reference
        }
    }
    std::cout << '-' << p_.nodes_[nn - 1]->name << std::endl;
    tmp = getSequenceFrom(p_.nodes_[nn - 1]->name);
    seq += reverse(substring(tmp, 0, p_.edges_[ne - 1]->q_end));
    std::reverse(seq.begin(), seq.end());
    out << seq << std::endl;
    return true;
}
