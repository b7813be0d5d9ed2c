// forwards to the nlohmann/json 3.11.3 single header shipped in this image (cudnn_frontend/thirdparty)
#pragma once
#include <cudnn_frontend/thirdparty/nlohmann/json.hpp>
