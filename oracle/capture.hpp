// TEST INFRASTRUCTURE (oracle/): shared by the reference-arm entry points.
//
// The reference solver reports its progress only on std::cout
// (/root/reference/src/LyapunovSolver.hpp:219-221,232-237).  These helpers
// capture that stream at 17 significant digits and parse the per-iteration
// residual estimates out of it, so a parity harness needs no solver change.
#ifndef RAILS_ORACLE_CAPTURE_HPP
#define RAILS_ORACLE_CAPTURE_HPP

#include <cstdlib>
#include <cstring>
#include <iostream>
#include <map>
#include <sstream>
#include <string>
#include <vector>

namespace rails_oracle
{

// Duck-typed ParameterList (same contract as the stub in
// /root/reference/test/LyapunovSolver_test.cpp:160-179).
class ParameterList
{
    std::map<std::string, double> params_;

public:
    ParameterList(int n, const char *const *keys, const double *vals)
    {
        for (int i = 0; i < n; ++i)
            params_[keys[i]] = vals[i];
    }

    template<typename T>
    T get(std::string const &name, T def)
    {
        auto it = params_.find(name);
        if (it == params_.end())
            return def;
        return (T)it->second;
    }
};

class CoutCapture
{
    std::ostringstream buf_;
    std::streambuf *old_;
    std::streamsize old_prec_;

public:
    CoutCapture()
    {
        buf_.precision(17);
        old_ = std::cout.rdbuf(buf_.rdbuf());
        old_prec_ = std::cout.precision(17);
    }
    ~CoutCapture()
    {
        std::cout.rdbuf(old_);
        std::cout.precision(old_prec_);
    }
    std::string str() const { return buf_.str(); }
};

// Parses "Iteration <i>. Estimate Lanczos, absolute: <a>, relative: <r>".
inline int parse_history(std::string const &log, std::vector<double> &abs_res,
                         std::vector<double> &rel_res)
{
    std::istringstream in(log);
    std::string line;
    while (std::getline(in, line))
    {
        if (line.compare(0, 10, "Iteration ") != 0)
            continue;
        size_t a = line.find("absolute: ");
        size_t r = line.find(", relative: ");
        if (a == std::string::npos || r == std::string::npos)
            continue;
        abs_res.push_back(std::strtod(line.c_str() + a + 10, nullptr));
        rel_res.push_back(std::strtod(line.c_str() + r + 12, nullptr));
    }
    return (int)abs_res.size();
}

}

#endif
