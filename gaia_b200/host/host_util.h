// Small helpers shared by the tet and cloth host adapters (folder creation, numbered file names, path parts).
#pragma once
#include <sys/stat.h>

#include <cstdio>
#include <string>

namespace gaia_b200 {

inline void createFolder(const std::string& p) {
    std::string cur;
    for (size_t i = 0; i <= p.size(); i++) {
        if (i == p.size() || p[i] == '/') {
            if (!cur.empty()) ::mkdir(cur.c_str(), 0755);
        }
        if (i < p.size()) cur.push_back(p[i]);
    }
}
inline std::string replaceAll(std::string s, const std::string& from, const std::string& to) {
    for (size_t at = s.find(from); at != std::string::npos; at = s.find(from, at + to.size())) s.replace(at, from.size(), to);
    return s;
}
inline std::string numbered(const char* prefix, int frameId, const char* ext, int width = 8) {
    char buf[64];
    std::snprintf(buf, sizeof(buf), "%s%0*d%s", prefix, width, frameId, ext);
    return buf;
}
inline std::string baseName(const std::string& path) {     // MF::IO::FileParts::name (file name without extension)
    const size_t s = path.find_last_of('/');
    std::string f = s == std::string::npos ? path : path.substr(s + 1);
    const size_t d = f.find_last_of('.');
    return d == std::string::npos ? f : f.substr(0, d);
}
inline std::string dirName(const std::string& path) {
    const size_t s = path.find_last_of('/');
    return s == std::string::npos ? std::string(".") : path.substr(0, s);
}

}  // namespace gaia_b200
