#include "groups.h"

#include <algorithm>
#include <cctype>
#include <cstdlib>
#include <fstream>
#include <iostream>
#include <sstream>

namespace speq
{
namespace
{
std::string without_blanks(std::string s)
{
    s.erase(std::remove_if(s.begin(), s.end(), [](char c) { return c == ' ' || c == '\t'; }), s.end());
    return s;
}

// same acceptance rule as the reference's is_integer: optional sign, then strtol must consume everything
bool looks_like_integer(std::string const & s)
{
    if (s.empty()) return false;
    if (!std::isdigit(static_cast<unsigned char>(s[0])) && s[0] != '-' && s[0] != '+') return false;
    char * end = nullptr;
    std::strtol(s.c_str(), &end, 10);
    return *end == '\0';
}

void mark(std::vector<int> & group_scaffolds, size_t first, size_t last, int variant)
{
    if (last + 1 > group_scaffolds.size()) group_scaffolds.resize(last + 1, -1);
    for (size_t i = first; i <= last; ++i) group_scaffolds[i] = variant;
}
}   // namespace

void file_to_map(std::filesystem::path const & path, std::vector<std::string> & group_names, std::vector<int> & group_scaffolds,
                 std::vector<int> & group_count)
{
    std::ifstream in(path);
    if (!in.is_open()) return;   // the reference silently yields empty vectors
    std::string line;
    int variant = -1;
    while (std::getline(in, line))
    {
        auto hash = line.find('#');
        if (hash != std::string::npos) line.erase(hash);
        auto colon = line.find(':');
        if (colon == std::string::npos) continue;
        auto open_p = line.find('('), close_p = line.find(')');
        group_names.push_back(line.substr(0, open_p));
        // std::stoi throws on a missing / non-numeric count exactly like the reference (src/file_to_map.cpp:42-43)
        group_count.push_back(std::stoi(line.substr(open_p + 1, close_p - open_p - 1)));
        ++variant;
        std::stringstream members(line.substr(colon + 1));
        std::string token;
        while (std::getline(members, token, ','))
        {
            auto dash = token.find('-');
            if (dash != std::string::npos)
            {
                std::string a = without_blanks(token.substr(0, dash)), b = without_blanks(token.substr(dash + 1));
                if (looks_like_integer(a) && looks_like_integer(b))
                    mark(group_scaffolds, static_cast<size_t>(std::atoi(a.c_str())), static_cast<size_t>(std::atoi(b.c_str())), variant);
                else
                    std::cerr << "Error in parsing groupings line: " << line << "\nA non-integer range was detected and ignored at: "
                              << without_blanks(token) << std::endl;
            }
            else
            {
                std::string a = without_blanks(token);
                if (looks_like_integer(a))
                {
                    size_t i = static_cast<size_t>(std::atoi(a.c_str()));
                    mark(group_scaffolds, i, i, variant);
                }
                else
                    std::cerr << "Error in parsing groupings line: " << line << "\nA non-integer index was detected and ignored at: "
                              << a << std::endl;
            }
        }
    }
}
}   // namespace speq
