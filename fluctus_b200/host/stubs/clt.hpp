#pragma once
#include "cl2.hpp"
namespace clt { class Kernel; }
