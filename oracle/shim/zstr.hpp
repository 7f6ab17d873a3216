#pragma once
#include <fstream>
namespace zstr { using ofstream = std::ofstream; }
