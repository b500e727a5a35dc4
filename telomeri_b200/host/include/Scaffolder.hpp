// Host mirror of the reference's Scaffolder (reference include/Scaffolder.hpp:11-34): FASTA/FASTQ
// loader and sequence stitching along the scaffold path.  With the two files main() loads (reads, contigs)
// write() hands the stitching to hera_scaffold (FASTA index + stitch kernel on the GPU) and checks that the
// library's chain of consensus paths is the Path it was given; otherwise the host restatement runs.
#ifndef TELOMERI_SCAFFOLDER_HPP
#define TELOMERI_SCAFFOLDER_HPP

#include "Path.hpp"

class Scaffolder {
public:
    const Path &p_;
    std::vector<std::string> names_;
    std::vector<std::string> sequences_;
    std::vector<std::string> files_;   // not in the reference: what load() was called with
    static unsigned long min_path_num_;   // what main() passed to constructConsensusPath (src/Main.cpp:344)

    explicit Scaffolder(const Path &p) : p_(p) {}

    bool load(const char *filename);
    bool load_host(const char *filename);   // the reference's loader (names_ / sequences_)
    /** characters start..end, both inclusive */
    std::string substring(std::string &str, ulong start, ulong end);
    std::string &reverse(std::string &);
    std::string reverse(std::string);
    std::string &getSequenceFrom(const std::string &name);
    bool write(const char *filename);
    void swapBases(std::string &str);
};

#endif
