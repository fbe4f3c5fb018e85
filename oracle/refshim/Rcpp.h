// TEST INFRASTRUCTURE — stand-in for <Rcpp.h>; see RcppArmadillo.h in this directory.
#pragma once
#include "RcppArmadillo.h"
