// ref_btf_fill.h — TEST INFRASTRUCTURE ONLY (see oracle/README.md).
// Fills the reference's own BTFBase (MLVQ/BTFBase.cuh) from an rtf_btf_desc table set the way
// CompressedBTF::ImportFromFolder does it (src/RayTracerFacility/MLVQ/CompressedBTF.cpp:152-191, 320-385), through the
// reference's own Init() methods.  Shared by ref_btf_driver.cpp (the evaluator alone) and ref_ie_driver.cpp (BTF
// materials inside the reference's closest-hit program).  The tables stay owned by the caller.
#pragma once
#include <cstring>

#include <BTFBase.cuh>

#include "../include/rtf/rtf.h"

static inline void ref_fill_btf(RayTracerFacility::BTFBase& b, const rtf_btf_desc* d, const float* beta_angles) {
  using namespace RayTracerFacility;
  b.m_hasData = true;
  b.m_materialOrder = 0, b.m_nColor = 0;                                   // CompressedBTF.cpp:93-94
  b.m_mPostScale = 1.0f;                                                   // :104
  b.m_useCosBeta = d->use_cos_beta != 0;                                   // :152
  b.m_numOfBeta = d->num_beta, b.m_numOfAlpha = d->num_alpha;              // :153-155
  b.m_numOfTheta = d->num_theta, b.m_numOfPhi = d->num_phi;                // :157-159
  b.m_stepAlpha = 180.f / (b.m_numOfAlpha - 1);                            // :189
  b.m_stepTheta = 90.0f / (b.m_numOfTheta - 1);                            // :190
  b.m_stepPhi = 360.0f / b.m_numOfPhi;                                     // :191
  SharedCoordinates& sc = b.m_sharedCoordinates;                           // SharedCoordinates::Set, :192-194
  memset((void*)&sc, 0, sizeof(sc));                                       // (m_phi etc. start at 0)
  sc.m_codeBtfFlag = d->code_btf_flag != 0;
  sc.m_useCosBeta = b.m_useCosBeta;
  sc.m_numOfBeta = b.m_numOfBeta;
  sc.m_numOfAlpha = b.m_numOfAlpha, sc.m_stepAlpha = b.m_stepAlpha;
  sc.m_numOfTheta = b.m_numOfTheta, sc.m_stepTheta = b.m_stepTheta;
  sc.m_numOfPhi = b.m_numOfPhi, sc.m_stepPhi = b.m_stepPhi;
  sc.m_deviceBetaAngles = const_cast<float*>(beta_angles);
  sc.m_hdrFlag = d->hdr != 0;                                              // :316-318
  b.m_hdr = d->hdr != 0, b.m_hdrValue = d->hdr_value;
  b.m_texCoordMultiplier = d->tex_coord_multiplier, b.m_gamma = d->gamma;
  auto& ab = b.m_pdf6.m_pdf4.m_pdf3.m_pdf2.m_iab.m_ab;                     // :320-332
  auto& iab = b.m_pdf6.m_pdf4.m_pdf3.m_pdf2.m_iab;
  auto& pdf1 = b.m_pdf6.m_pdf4.m_pdf3.m_pdf2.m_pdf1;
  auto& pdf2 = b.m_pdf6.m_pdf4.m_pdf3.m_pdf2;
  auto& pdf3 = b.m_pdf6.m_pdf4.m_pdf3;
  auto& pdf4 = b.m_pdf6.m_pdf4;
  pdf1.Init(b.m_numOfBeta);
  ab.Init();
  iab.Init(b.m_numOfBeta);
  pdf2.Init(b.m_numOfAlpha);
  pdf3.Init(b.m_numOfTheta);
  pdf4.Init(b.m_numOfPhi);
  b.m_pdf6.Init(d->pdf6_rows, d->pdf6_cols, d->rows_offset, d->cols_offset, b.m_nColor);
  // table shapes as ParseIntData / ParseFloatData return them through their (rows, cols) reference arguments
  // (:339-377: the column counts OVERWRITE what Init() put there — e.g. PDF2DColor::Init sets m_numOfAlpha to the beta
  // count), pointers as UploadDeviceData sets them
  pdf1.m_numOfBeta = d->num_beta, iab.m_numOfBeta = d->num_beta;
  pdf2.m_color.m_numOfAlpha = d->num_alpha, pdf2.m_luminance.m_numOfAlpha = d->num_alpha;
  pdf3.m_numOfTheta = d->num_theta, pdf4.m_numOfPhi = d->num_phi;
  b.m_pdf6.m_devicePdf6DSlices = const_cast<int*>(d->pdf6_slices);
  b.m_pdf6.m_devicePdf6DScales = const_cast<float*>(d->pdf6_scales);
  pdf1.m_numOfPdf1D = d->n_pdf1d, pdf1.m_devicePdf1DBasis = const_cast<float*>(d->pdf1d_basis);
  ab.m_numOfColors = d->n_colors, ab.m_numOfChannels = d->n_channels;
  ab.m_deviceVectorColorBasis = const_cast<float*>(d->colors);
  iab.m_numOfIndexSlices = d->n_index_slices, iab.m_deviceIndexAbBasis = const_cast<int*>(d->index_ab);
  pdf2.m_color.m_numOfPdf2D = d->n_pdf2d_color, pdf2.m_color.m_devicePdf2DColors = const_cast<int*>(d->pdf2d_colors);
  pdf2.m_luminance.m_numOfPdf2D = d->n_pdf2d_lum;
  pdf2.m_luminance.m_devicePdf2DSlices = const_cast<int*>(d->pdf2d_slices);
  pdf2.m_luminance.m_devicePdf2DScales = const_cast<float*>(d->pdf2d_scales);
  pdf2.m_numOfPdf2D = d->n_pdf2d, pdf2.m_lengthOfSlice = d->pdf2d_slice_len;
  pdf2.m_deviceIndexLuminanceColor = const_cast<int*>(d->pdf2d_indices);
  pdf3.m_numOfPdf3D = d->n_pdf3d;
  pdf3.m_devicePdf3DScales = const_cast<float*>(d->pdf3d_scales), pdf3.m_devicePdf3DSlices = const_cast<int*>(d->pdf3d_slices);
  pdf4.m_numOfPdf4D = d->n_pdf4d;
  pdf4.m_devicePdf4DScales = const_cast<float*>(d->pdf4d_scales), pdf4.m_devicePdf4DSlices = const_cast<int*>(d->pdf4d_slices);
  b.m_multiplier = d->multiplier;  // the loader sets hdr ? hdrValue : 1 (:382-386); rtf_btf_import_folder does the same
}
