// Shadow header: routes the reference's OpenCV include to the CPU oracle's mini-OpenCV (test infrastructure only).
#include "../../minicv.hpp"
