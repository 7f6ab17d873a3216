#pragma once
#include "upcxx_utils.hpp"
