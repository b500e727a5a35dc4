// Host mirror of the reference's PathWindow (reference include/PathWindow.hpp:8-36).  The device
// builds its windows inside group_pairs_kernel; this class exists for source compatibility and
// for printing (it is not used by main()).
#ifndef PATHWINDOW_HPP
#define PATHWINDOW_HPP

#include <map>

#include <Path.hpp>

class PathWindow {
private:
    std::vector<const Path *> piw_;
    std::map<ulong, int> frqs;
    int sum_frqs;
public:
    PathWindow(ulong l, ulong u, const std::vector<const Path *> &sp);
    std::pair<ulong, int> getLowestFrequencyEntry() const;
    std::pair<ulong, int> getHighestFrequencyEntry() const;
    int getSumFreqs() const { return sum_frqs; }
    friend std::ostream &operator<<(std::ostream &s, const PathWindow &pw);
};

#endif
