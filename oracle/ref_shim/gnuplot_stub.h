// Stand-in for the reference's vendored third-party gnuplot-iostream.h (needs Boost, absent
// here).  Pre-included with -include and -DGNUPLOT_IOSTREAM_H so that the reference's data.h
// (pulled in by derivatives_handler.h) compiles: data.h only specialises TextSender<>.
// Test infrastructure only; no plotting happens anywhere in this repo.
#pragma once
#include <ostream>
namespace gnuplotio {
template <typename T> struct TextSender {
    static void send(std::ostream &, const T &) {}
};
}
