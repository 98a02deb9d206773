# Build-time patch applied to the reference's src/heat3D.cpp ON THE FLY (sed | g++ -x c++ -);
# the patched text is never written to disk or committed.  Two insertions, both OUR code:
#
# (1) The reference leaves the "compute internal domain" slot between its GPU BEGIN / GPU END
#     banners empty (heat3D.cpp:606-612), so its interior never changes.  We fill that slot with
#     the triple loop over interior cells of the reference's own update expression
#     (heat3D.cu:128-131, identical to heat3D.cpp:663-666 with the halo operand replaced by the
#     in-domain neighbour); loop bounds as in its residual loop (heat3D.cpp:916-918).
/compute internal domain (no halos required)/a\
        for (unsigned i = 1; i < chunck[COORDINATE::X] - 1; ++i)\
            for (unsigned j = 1; j < chunck[COORDINATE::Y] - 1; ++j)\
                for (unsigned k = 1; k < chunck[COORDINATE::Z] - 1; ++k)\
                    T[i][j][k] = T0[i][j][k] +\
                        Dx * (T0[i + 1][j][k] - 2.0 * T0[i][j][k] + T0[i - 1][j][k]) +\
                        Dy * (T0[i][j + 1][k] - 2.0 * T0[i][j][k] + T0[i][j - 1][k]) +\
                        Dz * (T0[i][j][k + 1] - 2.0 * T0[i][j][k] + T0[i][j][k - 1]);
#
# (2) Test hooks placed after the error report and before the dump (heat3D.cpp:1005):
#     HEAT3D_REF_RAW=<path>  -> also write the final field as raw little-endian fp64, x fastest
#                               ((k*ny+j)*nx+i, the dump's own point order), for 1e-12 parity
#                               checks that the 6-digit text dump cannot support;
#     HEAT3D_REF_NODUMP=1    -> skip output/out.dat (61 B/point) when only the timing line matters.
/output the solution in a format readable by a post processor/i\
    if (const char* rawPath = std::getenv("HEAT3D_REF_RAW")) {\
        std::FILE* rawFile = std::fopen(rawPath, "wb");\
        if (rawFile) {\
            for (unsigned k = 0; k < chunck[COORDINATE::Z]; ++k)\
                for (unsigned j = 0; j < chunck[COORDINATE::Y]; ++j)\
                    for (unsigned i = 0; i < chunck[COORDINATE::X]; ++i)\
                        std::fwrite(&T[i][j][k], sizeof(double), 1, rawFile);\
            std::fclose(rawFile);\
        }\
    }\
    if (std::getenv("HEAT3D_REF_NODUMP")) { MPI_Finalize(); return 0; }
