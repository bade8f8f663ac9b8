#!/bin/bash
# Builds one kernel-tuning variant (hidden width 32 only) into ab/<name>.so:  tools/ab_build.sh <name> [-DFLAG=..]...
# See tools/ab_run.sh for the same-box comparison.
name=$1; shift
mkdir -p ab
make -s -C fv-srn_b200/csrc FAST=1 B=build_$name EXTRA="$*" TARGET=../../ab/$name.so -j8 2>&1 | grep -E "error|warning: v" ; ls -la ab/$name.so
