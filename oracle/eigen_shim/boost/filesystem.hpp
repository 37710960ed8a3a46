// Stand-in for the two boost::filesystem calls the reference's Load() makes
// (lib/GaussianProcess.cpp:196-237). TEST INFRASTRUCTURE ONLY (oracle build).
#ifndef GPR_ORACLE_BOOST_FS_SHIM
#define GPR_ORACLE_BOOST_FS_SHIM
#include <string>
#include <sys/stat.h>
namespace boost {
namespace filesystem {
class path {
public:
    path(const char* s) : s_(s) {}
    path(const std::string& s) : s_(s) {}
    const std::string& string() const { return s_; }
private:
    std::string s_;
};
inline bool exists(const path& p) {
    struct stat st;
    return ::stat(p.string().c_str(), &st) == 0;
}
inline bool is_directory(const path& p) {
    struct stat st;
    return ::stat(p.string().c_str(), &st) == 0 && S_ISDIR(st.st_mode);
}
}  // namespace filesystem
}  // namespace boost
#endif
