// TEST INFRASTRUCTURE ONLY -- see imgproc.hpp
#pragma once
#include "imgproc.hpp"
