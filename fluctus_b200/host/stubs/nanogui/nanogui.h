#pragma once
#include <functional>
namespace nanogui { class Screen; class Window; class Label; class ProgressBar; }
