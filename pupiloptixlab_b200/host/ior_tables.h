// ior_tables.h — named index-of-refraction data for the scene loader (physical constants: dielectric
// values from Hecht, Optics, 4th ed. at ~589 nm; conductor eta / k triples are the RGB fits
// distributed with Mitsuba).  Same names and values as the table the reference looks up
// (framework/render/material/ior.h:28-175), stored as text rows and parsed once.
#pragma once
#include <cstdlib>
#include <cstring>
#include <sstream>
#include <string>

namespace pupil::host {

// one row per material: name | ior
static const char kDielectricIorRows[] =
    "vacuum|1.0\n"
    "helium|1.000036\n"
    "hydrogen|1.000132\n"
    "air|1.000277\n"
    "carbon dioxide|1.00045\n"
    "water|1.3330\n"
    "acetone|1.36\n"
    "ethanol|1.361\n"
    "carbon tetrachloride|1.461\n"
    "glycerol|1.4729\n"
    "benzene|1.501\n"
    "silicone oil|1.52045\n"
    "bromine|1.661\n"
    "water ice|1.31\n"
    "fused quartz|1.458\n"
    "pyrex|1.470\n"
    "acrylic glass|1.49\n"
    "polypropylene|1.49\n"
    "bk7|1.5046\n"
    "sodium chloride|1.544\n"
    "amber|1.55\n"
    "pet|1.5750\n"
    "diamond|2.419\n";

// one row per material: name eta.r eta.g eta.b k.r k.g k.b
static const char kConductorIorRows[] =
    "a-C 2.93785 2.22242 1.96400 0.88555 0.79763 0.81356\n"
    "Ag 0.15494 0.11648 0.13809 4.81810 3.11562 2.14240\n"
    "Al 1.65394 0.87850 0.52012 9.20430 6.25621 4.82675\n"
    "Al2O3 2.12296 1.67556 1.61569 2.11325 1.66785 1.60811\n"
    "AlAs 3.59737 3.22615 2.21235 0.00067 -0.00050 0.00741\n"
    "AlAs_palik 3.59737 3.22615 2.21235 0.00067 -0.00050 0.00741\n"
    "AlSb 4.70830 4.05196 4.57419 -0.02918 0.09345 1.29784\n"
    "AlSb_palik 4.70830 4.05196 4.57419 -0.02918 0.09345 1.29784\n"
    "Au 0.14282 0.37414 1.43944 3.97472 2.38066 1.59981\n"
    "Be 4.17618 3.17830 2.77819 3.82730 3.00374 2.86293\n"
    "Be_palik 4.17618 3.17830 2.77819 3.82730 3.00374 2.86293\n"
    "Cr 4.36041 2.91052 1.65119 5.19538 4.22239 3.74700\n"
    "CsI 2.14035 1.69870 1.65890 0.00000 0.00000 0.00000\n"
    "CsI_palik 2.14035 1.69870 1.65890 0.00000 0.00000 0.00000\n"
    "Cu 0.19999 0.92209 1.09988 3.90464 2.44763 2.13765\n"
    "Cu_palik 0.21221 0.91805 1.10007 3.91324 2.45193 2.13213\n"
    "Cu2O 3.54426 2.94365 2.74619 0.11395 0.19341 0.60475\n"
    "Cu2O_palik 3.54426 2.94365 2.74619 0.11395 0.19341 0.60475\n"
    "CuO 3.24062 2.44261 2.20556 0.51999 0.56882 0.72064\n"
    "CuO_palik 3.24062 2.44261 2.20556 0.51999 0.56882 0.72064\n"
    "d-C 2.90008 2.29322 2.22275 0.00000 0.00000 0.00000\n"
    "d-C_palik 2.90008 2.29322 2.22275 0.00000 0.00000 0.00000\n"
    "Fe 2.76404 1.95417 1.62766 3.83077 2.73841 2.31812\n"
    "GaN 2.86583 2.29184 2.29848 0.00020 -0.00017 0.00128\n"
    "Ge 7.05832 4.78868 3.57814 0.50903 2.19196 2.05716\n"
    "Hg 2.39384 1.43697 0.90762 6.31420 4.36266 3.41454\n"
    "Hg_palik 2.39384 1.43697 0.90762 6.31420 4.36266 3.41454\n"
    "HgTe 4.76940 3.22413 2.65439 1.62853 1.57746 1.72592\n"
    "HgTe_palik 4.76940 3.22413 2.65439 1.62853 1.57746 1.72592\n"
    "Ir 3.07986 2.07777 1.61446 5.58028 4.05855 3.26033\n"
    "Ir_palik 3.07986 2.07777 1.61446 5.58028 4.05855 3.26033\n"
    "K 0.06391 0.04631 0.03810 2.09975 1.34607 0.91128\n"
    "K_palik 0.06391 0.04631 0.03810 2.09975 1.34607 0.91128\n"
    "Li 0.26525 0.19519 0.22045 3.53305 2.30618 1.66505\n"
    "Li_palik 0.26525 0.19519 0.22045 3.53305 2.30618 1.66505\n"
    "MgO 2.08521 1.64721 1.59150 0.00000 -0.00000 0.00000\n"
    "MgO_palik 2.08521 1.64721 1.59150 0.00000 -0.00000 0.00000\n"
    "Mo 4.47417 3.51799 2.77018 4.10240 3.41361 3.14393\n"
    "Mo_palik 4.47417 3.51799 2.77018 4.10240 3.41361 3.14393\n"
    "Na 0.06014 0.05602 0.06186 3.17254 2.10800 1.57575\n"
    "Na_palik 0.06014 0.05602 0.06186 3.17254 2.10800 1.57575\n"
    "Nb 3.41288 2.78427 2.39051 3.43408 2.73183 2.57445\n"
    "Nb_palik 3.41288 2.78427 2.39051 3.43408 2.73183 2.57445\n"
    "Ni 2.36225 1.65983 1.46395 4.48929 3.04369 2.34046\n"
    "Ni_palik 2.36225 1.65983 1.46395 4.48929 3.04369 2.34046\n"
    "Rh 2.58031 1.85624 1.55114 6.76790 4.69297 3.96766\n"
    "Rh_palik 2.58031 1.85624 1.55114 6.76790 4.69297 3.96766\n"
    "Se 4.08037 2.83343 2.81458 0.75378 0.63705 0.52047\n"
    "Se_palik 4.08037 2.83343 2.81458 0.75378 0.63705 0.52047\n"
    "SiC 3.16562 2.52061 2.47411 0.00000 -0.00000 0.00001\n"
    "SiC_palik 3.16562 2.52061 2.47411 0.00000 -0.00000 0.00001\n"
    "SnTe 4.51562 1.97692 1.27897 0.00000 0.00000 0.00000\n"
    "SnTe_palik 4.51562 1.97692 1.27897 0.00000 0.00000 0.00000\n"
    "Ta 2.05820 2.38802 2.62250 2.40293 1.73767 1.94291\n"
    "Ta_palik 2.05820 2.38802 2.62250 2.40293 1.73767 1.94291\n"
    "Te 7.37519 4.47257 2.63149 3.24919 3.51992 3.28521\n"
    "Te_palik 7.37519 4.47257 2.63149 3.24919 3.51992 3.28521\n"
    "ThF4 1.82684 1.43917 1.38471 0.00000 0.00000 0.00000\n"
    "ThF4_palik 1.82684 1.43917 1.38471 0.00000 0.00000 0.00000\n"
    "TiC 3.69261 2.83141 2.57683 3.25876 2.34656 2.16818\n"
    "TiC_palik 3.69261 2.83141 2.57683 3.25876 2.34656 2.16818\n"
    "TiN 1.64497 1.14800 1.37685 3.36132 1.93936 1.09967\n"
    "TiN_palik 1.64497 1.14800 1.37685 3.36132 1.93936 1.09967\n"
    "TiO2 3.44929 2.79576 2.89899 0.00010 -0.00009 0.00063\n"
    "TiO2_palik 3.44929 2.79576 2.89899 0.00010 -0.00009 0.00063\n"
    "VC 3.64981 2.74689 2.52731 3.06184 2.19400 1.95902\n"
    "VC_palik 3.64981 2.74689 2.52731 3.06184 2.19400 1.95902\n"
    "V 4.26843 3.50571 2.75528 3.48377 2.88322 3.10511\n"
    "V_palik 4.26843 3.50571 2.75528 3.48377 2.88322 3.10511\n"
    "VN 2.85952 2.11468 1.93597 3.02590 2.05174 1.61287\n"
    "VN_palik 2.85952 2.11468 1.93597 3.02590 2.05174 1.61287\n"
    "W 4.36142 3.29330 2.99191 3.49325 2.59934 2.26838\n"
    "none 0 0 0 1 1 1\n";

// LoadDielectricIor (ior.h:178-193): a number, else a name from the table, else the default
inline float dielectric_ior(const std::string& s, float def) {
    if (s.empty()) return def;
    char* end = nullptr;
    const float v = std::strtof(s.c_str(), &end);
    if (end && *end == '\0' && end != s.c_str()) return v;
    std::istringstream rows(kDielectricIorRows);
    std::string row;
    while (std::getline(rows, row)) {
        const size_t bar = row.find('|');
        if (bar != std::string::npos && row.compare(0, bar, s) == 0 && bar == s.size()) return std::strtof(row.c_str() + bar + 1, nullptr);
    }
    return def;
}

// LoadConductorIor (ior.h:196-207): false when the name is empty or unknown ("none" included)
inline bool conductor_ior(const std::string& name, float eta[3], float k[3]) {
    if (name.empty()) return false;
    std::istringstream rows(kConductorIorRows);
    std::string row;
    while (std::getline(rows, row)) {
        std::istringstream f(row);
        std::string n;
        f >> n;
        if (n != name) continue;
        f >> eta[0] >> eta[1] >> eta[2] >> k[0] >> k[1] >> k[2];
        return true;
    }
    return false;
}

}  // namespace pupil::host
