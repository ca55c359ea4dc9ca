# Out-of-tree registration, no change to the OpenCV source tree (samples/hal/README.md, CMakeLists.txt:956-958,
# :1069-1077):
#   make                                                   # builds opencv_b200/lib/libb200geom.so (sm_100a)
#   cmake -S <opencv> -B <build> -DOpenCV_HAL_DIR=<this repo>/hal/cuda_b200 ...
get_filename_component(_b200_root "${CMAKE_CURRENT_LIST_DIR}/../.." ABSOLUTE)
set(OpenCV_HAL_FOUND TRUE)
set(OpenCV_HAL_VERSION "0.1.0")
set(OpenCV_HAL_LIBRARIES "${_b200_root}/opencv_b200/lib/libb200geom.so")
set(OpenCV_HAL_HEADERS "cuda_b200_hal.hpp")
set(OpenCV_HAL_INCLUDE_DIRS "${CMAKE_CURRENT_LIST_DIR}/include" "${_b200_root}/include")
