/* Empty stand-in so that the reference's lib/knn/src/cpu/vision.h (which includes
 * <torch/extension.h> but uses nothing from it) compiles with plain g++ in a second.
 * Used only by oracle/Makefile when building oracle/_ref from /root/reference. */
#pragma once
