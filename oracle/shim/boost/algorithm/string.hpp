// ORACLE build shim (test infrastructure): the reference's src/Gas_optics_rrtmgp.cpp:29 includes
// <boost/algorithm/string.hpp> for a single boost::trim call (:86). Boost is absent here; this
// header supplies that one function so the reference source compiles unmodified.
#pragma once
#include <string>
namespace boost
{
    inline void trim(std::string& s)
    {
        const char* ws = " \t\n\r\f\v";
        const size_t b = s.find_first_not_of(ws);
        if (b == std::string::npos) { s.clear(); return; }
        const size_t e = s.find_last_not_of(ws);
        s = s.substr(b, e - b + 1);
    }
}
