// Shim for building /root/reference/lzf.cpp alone (oracle/Makefile target "ref"): the reference's own
// stdafx.h pulls Boost/OpenCV/PROJ/trimesh2, none of which lzf.cpp uses.
#pragma once
#include <stdio.h>
#include <iostream>
