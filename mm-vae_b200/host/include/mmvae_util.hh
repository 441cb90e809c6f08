// Host-side helpers of the B200 build: logging / check macros with the reference's names and
// behaviour (include/utils/util.hh:35-107 — TLOG/ELOG/CHK/ASSERT/ERR_RET print to stderr and the
// program returns EXIT_FAILURE), small string utilities used by the option parsers.
#ifndef MMVAE_B200_UTIL_HH_
#define MMVAE_B200_UTIL_HH_

#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <ctime>
#include <iostream>
#include <sstream>
#include <string>
#include <sys/stat.h>
#include <vector>

inline std::string
mmv_curr_time()
{
    char buf[64];
    std::time_t t = std::time(nullptr);
    std::strftime(buf, sizeof(buf), "%c", std::localtime(&t));
    return std::string(buf);
}

#define TLOG(msg) do { std::cerr << "[" << mmv_curr_time() << "] " << msg << std::endl; } while (0)
#define ELOG(msg) do { std::cerr << "[Error] [" << mmv_curr_time() << "] " << msg << std::endl; } while (0)
#define WLOG(msg) do { std::cerr << "[Warning] [" << mmv_curr_time() << "] " << msg << std::endl; } while (0)
#define CHK(cond) do { if ((cond) != EXIT_SUCCESS) return EXIT_FAILURE; } while (0)
#define ERR_RET(cond, msg) do { if (cond) { ELOG(msg); return EXIT_FAILURE; } } while (0)
#define ASSERT(cond, msg) do { if (!(cond)) { ELOG(msg); std::exit(1); } } while (0)
// C-ABI call: nonzero -> log mmv_last_error() and fail like the reference's CHK
#define MMV_OK(call) do { if ((call) != 0) { ELOG(#call << ": " << mmv_last_error()); return EXIT_FAILURE; } } while (0)
#define MMV_OK_EXIT(call) do { if ((call) != 0) { ELOG(#call << ": " << mmv_last_error()); std::exit(1); } } while (0)

inline bool
file_exists(const std::string &f)
{
    if (f.empty()) return false;
    struct stat st;
    return stat(f.c_str(), &st) == 0;
}

// "10,20,30" -> {10, 20, 30}  (std_util.hh copy_int_arr semantics: comma separated integers)
inline void
copy_int_arr(const std::string &s, std::vector<int64_t> &out)
{
    out.clear();
    std::stringstream ss(s);
    std::string tok;
    while (std::getline(ss, tok, ','))
        if (!tok.empty()) out.push_back(std::stol(tok));
}

// util.hh:97-107
inline std::string
zeropad(const int64_t t, const int64_t tmax)
{
    std::string tt = std::to_string(t), mm = std::to_string(tmax);
    return std::string(mm.size() > tt.size() ? mm.size() - tt.size() : 0, '0') + tt;
}

#endif
