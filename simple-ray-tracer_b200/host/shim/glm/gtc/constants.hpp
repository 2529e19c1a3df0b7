// part of the minimal glm stand-in: everything lives in glm.hpp
#pragma once
#include "../glm.hpp"
