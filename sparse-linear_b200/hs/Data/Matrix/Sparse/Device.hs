{-# LANGUAGE FlexibleContexts #-}
{-# LANGUAGE ForeignFunctionInterface #-}
{-# LANGUAGE RecordWildCards #-}
{-# LANGUAGE ScopedTypeVariables #-}

-- | Data.Matrix.Sparse.Device -- the reference-side binding of libsparselinear_b200.so.
--
-- UNVERIFIED SOURCE: this image has no GHC, so this module has never been compiled.  It is the
-- shim a maintainer would add to the sparse-linear package (see INTEGRATION.md); it follows the
-- conventions of the package's existing bindings:
--
--   * matrix arguments are passed exactly as 'withConstMatrix' lends them
--     (Data/Matrix/Sparse/Foreign.hs:24-41): CInt nrows, CInt ncols, Ptr CInt colptr,
--     Ptr CInt rowidx, Ptr a values;
--   * one class with an instance per element type selects the d/z symbol, like 'Umfpack'
--     (suitesparse/src/Numeric/LinearAlgebra/Umfpack/Internal.hs:61-67,117-162);
--   * the device matrix is an opaque pointer in a 'ForeignPtr' whose finalizer is imported by
--     address (Umfpack/Internal.hs:110-113);
--   * a negative status becomes 'errorWithStackTrace' with the library's message, which already
--     carries the reference's prefix ("axpy_: ...", "mm: ...", "compress: ...").
--
-- The functions exported here have the types of their namesakes in Data.Matrix.Sparse
-- (Data/Matrix/Sparse.hs:13-30), so replacing the import is the whole integration:
--
-- > import Data.Matrix.Sparse hiding (mulV, axpy_, axpy, transpose, ctrans, hermitian, compress, fromTriples, lin)
-- > import Data.Matrix.Sparse.Device
module Data.Matrix.Sparse.Device
       ( DeviceMatrix, SlcElem
       , toDevice, fromDevice
       , mulV, axpy, axpy_, mulM_, mulVD
       , mm, transpose, ctrans, hermitian
       , compress, fromTriples, lin
       ) where

import Control.Monad (when)
import Data.Complex (Complex)
import qualified Data.Vector.Storable as VS
import qualified Data.Vector.Storable.Mutable as VSM
import qualified Data.Vector.Unboxed as U
import Data.Vector.Unboxed (Unbox, Vector)
import Foreign.C.String (CString, peekCString)
import Foreign.C.Types (CDouble, CInt(..), CLLong(..))
import Foreign.ForeignPtr (ForeignPtr, newForeignPtr, withForeignPtr)
import Foreign.Marshal.Alloc (alloca)
import Foreign.Marshal.Utils (with)
import Foreign.Ptr (FunPtr, Ptr, castPtr, nullPtr)
import Foreign.Storable (Storable, peek)
import GHC.Stack (errorWithStackTrace)
import System.IO.Unsafe (unsafePerformIO)

import Data.Matrix.Sparse (Matrix(..))
import Data.Matrix.Sparse.Foreign (fromForeign, withConstMatrix)

-- | Opaque slc_matrix (include/slc.h).
data SlcMatrix

-- | A matrix resident in HBM: uploaded once, multiplied many times.
newtype DeviceMatrix a = DeviceMatrix (ForeignPtr SlcMatrix)

foreign import ccall unsafe "slc_last_error" slc_last_error :: IO CString
foreign import ccall unsafe "&slc_free" slc_free :: FunPtr (Ptr SlcMatrix -> IO ())
foreign import ccall unsafe "slc_dims"
  slc_dims :: Ptr SlcMatrix -> Ptr CInt -> Ptr CInt -> Ptr CLLong -> IO CInt

type Upload a = CInt -> CInt -> Ptr CInt -> Ptr CInt -> Ptr a -> Ptr (Ptr SlcMatrix) -> IO CInt
type Download a = Ptr SlcMatrix -> Ptr CInt -> Ptr CInt -> Ptr a -> IO CInt
type MulV a = Ptr SlcMatrix -> Ptr a -> Ptr a -> IO CInt
type MulM a = Ptr SlcMatrix -> Ptr a -> CLLong -> Ptr a -> CLLong -> CInt -> IO CInt
type Binary = Ptr SlcMatrix -> Ptr SlcMatrix -> Ptr (Ptr SlcMatrix) -> IO CInt
type Transpose = Ptr SlcMatrix -> CInt -> Ptr (Ptr SlcMatrix) -> IO CInt
type Hermitian = Ptr SlcMatrix -> Ptr CInt -> IO CInt
type Compress a = CInt -> CInt -> CLLong -> Ptr CInt -> Ptr CInt -> Ptr a
               -> Ptr (Ptr SlcMatrix) -> Ptr CLLong -> IO CInt
type Lin a = Ptr a -> Ptr SlcMatrix -> Ptr a -> Ptr SlcMatrix -> Ptr (Ptr SlcMatrix) -> IO CInt

foreign import ccall "slc_d_upload" slc_d_upload :: Upload Double
foreign import ccall "slc_z_upload" slc_z_upload :: Upload (Complex Double)
foreign import ccall "slc_d_download" slc_d_download :: Download Double
foreign import ccall "slc_z_download" slc_z_download :: Download (Complex Double)
foreign import ccall "slc_d_mulv" slc_d_mulv :: MulV Double
foreign import ccall "slc_z_mulv" slc_z_mulv :: MulV (Complex Double)
foreign import ccall "slc_d_axpy" slc_d_axpy :: MulV Double
foreign import ccall "slc_z_axpy" slc_z_axpy :: MulV (Complex Double)
foreign import ccall "slc_d_mulm" slc_d_mulm :: MulM Double
foreign import ccall "slc_z_mulm" slc_z_mulm :: MulM (Complex Double)
foreign import ccall "slc_d_mm" slc_d_mm :: Binary
foreign import ccall "slc_z_mm" slc_z_mm :: Binary
foreign import ccall "slc_d_transpose" slc_d_transpose :: Transpose
foreign import ccall "slc_z_transpose" slc_z_transpose :: Transpose
foreign import ccall "slc_d_hermitian" slc_d_hermitian :: Hermitian
foreign import ccall "slc_z_hermitian" slc_z_hermitian :: Hermitian
foreign import ccall "slc_d_compress" slc_d_compress :: Compress Double
foreign import ccall "slc_z_compress" slc_z_compress :: Compress (Complex Double)
foreign import ccall "slc_d_lin" slc_d_lin :: Lin Double
foreign import ccall "slc_z_lin" slc_z_lin :: Lin (Complex Double)

-- | Element types of the path: Double and Complex Double
-- (Data/Complex/Enhanced.hs:35-49).  Complex values cross the boundary interleaved (re, im),
-- which is Storable's layout for 'Complex Double'.
class (Num a, Storable a, Unbox a) => SlcElem a where
  c_upload :: Upload a
  c_download :: Download a
  c_mulv :: MulV a
  c_axpy :: MulV a
  c_mulm :: MulM a
  c_mm :: proxy a -> Binary
  c_transpose :: proxy a -> Transpose
  c_hermitian :: proxy a -> Hermitian
  c_compress :: Compress a
  c_lin :: Lin a

instance SlcElem Double where
  c_upload = slc_d_upload
  c_download = slc_d_download
  c_mulv = slc_d_mulv
  c_axpy = slc_d_axpy
  c_mulm = slc_d_mulm
  c_mm _ = slc_d_mm
  c_transpose _ = slc_d_transpose
  c_hermitian _ = slc_d_hermitian
  c_compress = slc_d_compress
  c_lin = slc_d_lin

instance SlcElem (Complex Double) where
  c_upload = slc_z_upload
  c_download = slc_z_download
  c_mulv = slc_z_mulv
  c_axpy = slc_z_axpy
  c_mulm = slc_z_mulm
  c_mm _ = slc_z_mm
  c_transpose _ = slc_z_transpose
  c_hermitian _ = slc_z_hermitian
  c_compress = slc_z_compress
  c_lin = slc_z_lin

-- | Status check: negative is fatal, as in Umfpack.hs:64-67.
check :: CInt -> IO ()
check rc = when (rc < 0) $ do
  msg <- slc_last_error >>= peekCString
  errorWithStackTrace msg

adopt :: Ptr (Ptr SlcMatrix) -> IO (DeviceMatrix a)
adopt out = peek out >>= fmap DeviceMatrix . newForeignPtr slc_free

-- | Upload a host matrix (the arrays are lent for the call only, Foreign.hs:33-37).
toDevice :: SlcElem a => Matrix Vector a -> IO (DeviceMatrix a)
toDevice mat =
  withConstMatrix mat $ \nr nc ptrs rows vals ->
  alloca $ \out -> do
    c_upload nr nc ptrs rows vals out >>= check
    adopt out

-- | Copy a device matrix back into a host record ('fromForeign' with copy = True owns the
-- buffers it is given, Foreign.hs:43-55; here they are vectors this function allocates).
fromDevice :: forall a. SlcElem a => DeviceMatrix a -> IO (Matrix Vector a)
fromDevice (DeviceMatrix fp) =
  withForeignPtr fp $ \p ->
  alloca $ \pnr -> alloca $ \pnc -> alloca $ \pnz -> do
    slc_dims p pnr pnc pnz >>= check
    nr <- peek pnr
    nc <- peek pnc
    nz <- fromIntegral <$> peek pnz
    ptrs <- VSM.new (fromIntegral nc + 1)
    rows <- VSM.new nz
    vals <- VSM.new nz :: IO (VSM.IOVector a)
    VSM.unsafeWith ptrs $ \pp -> VSM.unsafeWith rows $ \pr -> VSM.unsafeWith vals $ \pv -> do
      c_download p pp pr pv >>= check
      fromForeign True nr nc pp pr pv

onDevice :: SlcElem a => Matrix Vector a -> (Ptr SlcMatrix -> IO b) -> IO b
onDevice mat act = do
  DeviceMatrix fp <- toDevice mat
  withForeignPtr fp act

-- | y <- A x + y in place (axpy_, Data/Matrix/Sparse.hs:433-453), operands in host memory.
-- Dimension errors are raised here, before the call, with the reference's messages.
axpy_ :: SlcElem a => Matrix Vector a -> VSM.IOVector a -> VSM.IOVector a -> IO ()
axpy_ mat@Matrix{..} xs ys
  | VSM.length xs /= ncols = oops ("column dimension " ++ show ncols
                                   ++ " does not match operand dimension " ++ show (VSM.length xs))
  | VSM.length ys /= nrows = oops ("row dimension " ++ show nrows
                                   ++ " does not match result dimension " ++ show (VSM.length ys))
  | otherwise =
      onDevice mat $ \p ->
      VSM.unsafeWith xs $ \px -> VSM.unsafeWith ys $ \py -> c_axpy p px py >>= check
  where oops str = errorWithStackTrace ("axpy_: " ++ str)

axpy :: SlcElem a => Matrix Vector a -> Vector a -> Vector a -> Vector a
axpy a x y = unsafePerformIO $ do
  xs <- VS.thaw (VS.convert x)
  ys <- VS.thaw (VS.convert y)
  axpy_ a xs ys
  VS.convert <$> VS.freeze ys

-- | A x (mulV, Data/Matrix/Sparse.hs:464-471).
mulV :: SlcElem a => Matrix Vector a -> Vector a -> Vector a
mulV a x = unsafePerformIO $ do
  d <- toDevice a
  mulVD d (nrows a) (ncols a) x

-- | A x for a matrix that is already resident: repeated products never re-copy A.
mulVD :: SlcElem a => DeviceMatrix a -> Int -> Int -> Vector a -> IO (Vector a)
mulVD (DeviceMatrix fp) nr nc x
  | U.length x /= nc = errorWithStackTrace ("axpy_: column dimension " ++ show nc
                                            ++ " does not match operand dimension " ++ show (U.length x))
  | otherwise = withForeignPtr fp $ \p -> do
      ys <- VSM.new nr
      VS.unsafeWith (VS.convert x) $ \px -> VSM.unsafeWith ys $ \py -> c_mulv p px py >>= check
      VS.convert <$> VS.freeze ys

-- | Y[:, j] <- A X[:, j] for k columns of column-major blocks with leading dimension n: the whole
-- of FEAST's multiplyWork (feast/src/Numeric/LinearAlgebra/Feast.hs:203-208) in one call.
mulM_ :: SlcElem a => DeviceMatrix a -> Int -> Int -> VSM.IOVector a -> VSM.IOVector a -> IO ()
mulM_ (DeviceMatrix fp) n k xs ys =
  withForeignPtr fp $ \p ->
  VSM.unsafeWith xs $ \px -> VSM.unsafeWith ys $ \py ->
    c_mulm p px (fromIntegral n) py (fromIntegral n) (fromIntegral k) >>= check

binary :: SlcElem a => Binary -> Matrix Vector a -> Matrix Vector a -> Matrix Vector a
binary f a b = unsafePerformIO $
  onDevice a $ \pa -> onDevice b $ \pb ->
  alloca $ \out -> do
    f pa pb out >>= check
    adopt out >>= fromDevice

-- | A B (mm, Data/Matrix/Sparse.hs:691-702).
mm :: forall a. SlcElem a => Matrix Vector a -> Matrix Vector a -> Matrix Vector a
mm a b
  | ncols a /= nrows b = errorWithStackTrace "mm: inner dimension mismatch"
  | otherwise = binary (c_mm ([] :: [a])) a b

transposeWith :: forall a. SlcElem a => CInt -> Matrix Vector a -> Matrix Vector a
transposeWith conj a = unsafePerformIO $
  onDevice a $ \pa -> alloca $ \out -> do
    c_transpose ([] :: [a]) pa conj out >>= check
    adopt out >>= fromDevice

transpose, ctrans :: SlcElem a => Matrix Vector a -> Matrix Vector a
transpose = transposeWith 0   -- Data/Matrix/Sparse.hs:301-329
ctrans = transposeWith 1      -- Data/Matrix/Sparse.hs:371-375

-- | ctrans m == m (Data/Matrix/Sparse.hs:377-379), decided on the device.
hermitian :: forall a. SlcElem a => Matrix Vector a -> Bool
hermitian a = unsafePerformIO $
  onDevice a $ \pa -> alloca $ \res -> do
    c_hermitian ([] :: [a]) pa res >>= check
    (/= 0) <$> peek res

-- | COO -> CSC (compress, Data/Matrix/Sparse.hs:184-255).  The library reports the first offending
-- position with the reference's wording ("compress: row index out of bounds (0,n) at i").
compress :: SlcElem a => Int -> Int -> Vector Int -> Vector Int -> Vector a -> Matrix Vector a
compress nr nc rows cols vals
  | U.length rows /= U.length cols = errorWithStackTrace "compress: row and column array lengths differ"
  | U.length rows /= U.length vals = errorWithStackTrace "compress: row and value array lengths differ"
  | otherwise = unsafePerformIO $
      VS.unsafeWith (VS.map fromIntegral (VS.convert rows)) $ \pr ->
      VS.unsafeWith (VS.map fromIntegral (VS.convert cols)) $ \pc ->
      VS.unsafeWith (VS.convert vals) $ \pv ->
      alloca $ \out -> alloca $ \bad -> do
        c_compress (fromIntegral nr) (fromIntegral nc) (fromIntegral (U.length rows)) pr pc pv out bad
          >>= check
        adopt out >>= fromDevice

fromTriples :: SlcElem a => Int -> Int -> [(Int, Int, a)] -> Matrix Vector a
fromTriples nr nc triples =
  let (rows, cols, vals) = unzip3 triples
  in compress nr nc (U.fromList rows) (U.fromList cols) (U.fromList vals)

-- | alpha A + beta B on the union pattern (lin, Data/Matrix/Sparse.hs:426-431).
lin :: SlcElem a => a -> Matrix Vector a -> a -> Matrix Vector a -> Matrix Vector a
lin alpha a beta b
  | nrows a /= nrows b = errorWithStackTrace "glin: row number mismatch"
  | ncols a /= ncols b = errorWithStackTrace "glin: column number mismatch"
  | otherwise = unsafePerformIO $
      onDevice a $ \pa -> onDevice b $ \pb ->
      with alpha $ \pal -> with beta $ \pbe ->
      alloca $ \out -> do
        c_lin pal pa pbe pb out >>= check
        adopt out >>= fromDevice
