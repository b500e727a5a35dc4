// Host mirror of the reference's Scaffolder (reference include/Scaffolder.hpp:11-34): FASTA/FASTQ
// loader and sequence stitching along the scaffold path.  Host code.
#ifndef TELOMERI_SCAFFOLDER_HPP
#define TELOMERI_SCAFFOLDER_HPP

#include "Path.hpp"

class Scaffolder {
public:
    const Path &p_;
    std::vector<std::string> names_;
    std::vector<std::string> sequences_;

    explicit Scaffolder(const Path &p) : p_(p) {}

    bool load(const char *filename);
    /** characters start..end, both inclusive */
    std::string substring(std::string &str, ulong start, ulong end);
    std::string &reverse(std::string &);
    std::string reverse(std::string);
    std::string &getSequenceFrom(const std::string &name);
    bool write(const char *filename);
    void swapBases(std::string &str);
};

#endif
