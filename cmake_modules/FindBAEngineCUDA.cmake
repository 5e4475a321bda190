# FindBAEngineCUDA.cmake — locates the B200 bundle-adjustment engine (sits next to the reference's cmake_modules/FindG2O.cmake,
# which it replaces for the src/adjuster path: the reference finds g2o at CMakeLists.txt:28-30 but never links it).
#
#   find_package(BAEngineCUDA REQUIRED)
#   target_link_libraries(<target> PRIVATE BAEngineCUDA::ba_engine)
#
# Sets BAEngineCUDA_FOUND, BAEngineCUDA_INCLUDE_DIR (ba_engine.h), BAEngineCUDA_ADAPTER_DIR (host/adjuster/BundleAdjuster.h),
# BAEngineCUDA_LIBRARY (libba_engine.so, sm_100a only).  Hint with -DBAENGINE_ROOT=<this repo or its install prefix>.
find_path(BAEngineCUDA_INCLUDE_DIR NAMES ba_engine.h
          HINTS ${BAENGINE_ROOT} $ENV{BAENGINE_ROOT} PATH_SUFFIXES include)
find_library(BAEngineCUDA_LIBRARY NAMES ba_engine
             HINTS ${BAENGINE_ROOT} $ENV{BAENGINE_ROOT} PATH_SUFFIXES lib 3dreconstruction_b200/csrc build)
find_path(BAEngineCUDA_ADAPTER_DIR NAMES adjuster/BundleAdjuster.h
          HINTS ${BAENGINE_ROOT} $ENV{BAENGINE_ROOT} PATH_SUFFIXES 3dreconstruction_b200/host include/b200ba)

include(FindPackageHandleStandardArgs)
find_package_handle_standard_args(BAEngineCUDA DEFAULT_MSG BAEngineCUDA_LIBRARY BAEngineCUDA_INCLUDE_DIR)

if(BAEngineCUDA_FOUND AND NOT TARGET BAEngineCUDA::ba_engine)
  add_library(BAEngineCUDA::ba_engine SHARED IMPORTED)
  set_target_properties(BAEngineCUDA::ba_engine PROPERTIES
    IMPORTED_LOCATION "${BAEngineCUDA_LIBRARY}"
    INTERFACE_INCLUDE_DIRECTORIES "${BAEngineCUDA_INCLUDE_DIR}")
endif()
mark_as_advanced(BAEngineCUDA_INCLUDE_DIR BAEngineCUDA_LIBRARY BAEngineCUDA_ADAPTER_DIR)
