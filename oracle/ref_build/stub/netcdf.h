// Stub netcdf.h: declarations only, so that the reference's YAKL_netcdf.h (pulled in by
// dynamics/awfl/awfl_const.h:6) parses when building the oracle harness. Nothing here is ever
// called on the AWFL path, so no NetCDF library is linked. Signatures follow the public NetCDF-C API.
#pragma once
#include <stddef.h>
typedef int nc_type;
#define NC_NOERR 0
#define NC_BYTE 1
#define NC_CHAR 2
#define NC_SHORT 3
#define NC_INT 4
#define NC_FLOAT 5
#define NC_DOUBLE 6
#define NC_UBYTE 7
#define NC_USHORT 8
#define NC_UINT 9
#define NC_INT64 10
#define NC_UINT64 11
#define NC_NOWRITE 0x0000
#define NC_WRITE 0x0001
#define NC_CLOBBER 0x0000
#define NC_NOCLOBBER 0x0004
#define NC_NETCDF4 0x1000
#define NC_UNLIMITED 0L
#define NC_MAX_NAME 256
#define NC_MAX_VAR_DIMS 1024
#ifdef __cplusplus
extern "C" {
#endif
const char *nc_strerror(int ncerr);
int nc_open(const char *path, int mode, int *ncidp);
int nc_create(const char *path, int cmode, int *ncidp);
int nc_close(int ncid);
int nc_def_dim(int ncid, const char *name, size_t len, int *idp);
int nc_def_var(int ncid, const char *name, nc_type xtype, int ndims, const int *dimidsp, int *varidp);
int nc_inq_dim(int ncid, int dimid, char *name, size_t *lenp);
int nc_inq_dimid(int ncid, const char *name, int *idp);
int nc_inq_unlimdim(int ncid, int *unlimdimidp);
int nc_inq_var(int ncid, int varid, char *name, nc_type *xtypep, int *ndimsp, int *dimidsp, int *nattsp);
int nc_inq_varid(int ncid, const char *name, int *varidp);
int nc_get_var_double(int ncid, int varid, double *ip);
int nc_put_var_double(int ncid, int varid, const double *op);
int nc_put_vara_double(int ncid, int varid, const size_t *startp, const size_t *countp, const double *op);
int nc_get_var_float(int ncid, int varid, float *ip);
int nc_put_var_float(int ncid, int varid, const float *op);
int nc_put_vara_float(int ncid, int varid, const size_t *startp, const size_t *countp, const float *op);
int nc_get_var_int(int ncid, int varid, int *ip);
int nc_put_var_int(int ncid, int varid, const int *op);
int nc_put_vara_int(int ncid, int varid, const size_t *startp, const size_t *countp, const int *op);
int nc_get_var_long(int ncid, int varid, long *ip);
int nc_put_var_long(int ncid, int varid, const long *op);
int nc_put_vara_long(int ncid, int varid, const size_t *startp, const size_t *countp, const long *op);
int nc_get_var_longlong(int ncid, int varid, long long *ip);
int nc_put_var_longlong(int ncid, int varid, const long long *op);
int nc_put_vara_longlong(int ncid, int varid, const size_t *startp, const size_t *countp, const long long *op);
int nc_get_var_schar(int ncid, int varid, signed char *ip);
int nc_put_var_schar(int ncid, int varid, const signed char *op);
int nc_put_vara_schar(int ncid, int varid, const size_t *startp, const size_t *countp, const signed char *op);
int nc_get_var_short(int ncid, int varid, short *ip);
int nc_put_var_short(int ncid, int varid, const short *op);
int nc_put_vara_short(int ncid, int varid, const size_t *startp, const size_t *countp, const short *op);
int nc_get_var_text(int ncid, int varid, char *ip);
int nc_put_var_text(int ncid, int varid, const char *op);
int nc_put_vara_text(int ncid, int varid, const size_t *startp, const size_t *countp, const char *op);
int nc_get_var_uchar(int ncid, int varid, unsigned char *ip);
int nc_put_var_uchar(int ncid, int varid, const unsigned char *op);
int nc_put_vara_uchar(int ncid, int varid, const size_t *startp, const size_t *countp, const unsigned char *op);
int nc_get_var_uint(int ncid, int varid, unsigned int *ip);
int nc_put_var_uint(int ncid, int varid, const unsigned int *op);
int nc_put_vara_uint(int ncid, int varid, const size_t *startp, const size_t *countp, const unsigned int *op);
int nc_get_var_ulonglong(int ncid, int varid, unsigned long long *ip);
int nc_put_var_ulonglong(int ncid, int varid, const unsigned long long *op);
int nc_put_vara_ulonglong(int ncid, int varid, const size_t *startp, const size_t *countp, const unsigned long long *op);
int nc_get_var_ushort(int ncid, int varid, unsigned short *ip);
int nc_put_var_ushort(int ncid, int varid, const unsigned short *op);
int nc_put_vara_ushort(int ncid, int varid, const size_t *startp, const size_t *countp, const unsigned short *op);
#ifdef __cplusplus
}
#endif
