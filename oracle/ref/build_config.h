/* What cmake would generate from the reference's config/build_config.h.in
 * (values: CMakeLists.txt:3).  Written by hand: the reference's own build
 * system is not run (it needs GLFW/assimp/asio). */
#define BUILD_PROJECT_NAME "INTERCESSION"
#define BUILD_VERSION_MAJOR 0
#define BUILD_VERSION_MINOR 12
#define BUILD_VERSION_PATCH 0
