// groups.h -- the groups file: which scaffolds (by FASTA order) belong to which variant.
// Grammar and side effects follow the reference's speq::file_to_map (src/file_to_map.cpp:20-119); SURVEY.md section 8b.
#pragma once
#include <filesystem>
#include <string>
#include <vector>

namespace speq
{
// group_names[v]; group_scaffolds[scaffold] = v (-1 where no line mentions the scaffold); group_count[v] = the "(count)"
void file_to_map(std::filesystem::path const & path, std::vector<std::string> & group_names, std::vector<int> & group_scaffolds,
                 std::vector<int> & group_count);
}   // namespace speq
