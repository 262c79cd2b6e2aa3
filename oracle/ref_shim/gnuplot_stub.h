// Stand-in for the reference's vendored third-party gnuplot-iostream.h (needs Boost, absent
// here).  Pre-included with -include and -DGNUPLOT_IOSTREAM_H so that the reference's data.h
// (pulled in by derivatives_handler.h) compiles: data.h only specialises TextSender<>.
// Test infrastructure only; no plotting happens anywhere in this repo.
#pragma once
#include <ostream>
#include <unistd.h>
namespace gnuplotio {
template <typename T> struct TextSender {
    static void send(std::ostream &, const T &) {}
};
}
// diagnostics.cpp drives gnuplot through this class; here every command is dropped (no plotting).
class Gnuplot {
public:
    Gnuplot() {}
    template <typename T> Gnuplot &operator<<(const T &) { return *this; }
    Gnuplot &operator<<(std::ostream &(*)(std::ostream &)) { return *this; }
    template <typename T> Gnuplot &send1d(const T &) { return *this; }
    template <typename T> std::string file1d(const T &) { return std::string(); }
    template <typename T> std::string file1d(const T &, const std::string &) { return std::string(); }
};
