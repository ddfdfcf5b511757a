{-# LANGUAGE ExistentialQuantification #-}
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
-- UNCHANGED CALLERS.  `mulV a x` on an ordinary host 'Matrix' must not re-upload A on every call
-- (SURVEY.md section 8(b): "a memo keyed on the host matrix").  'resident' below keeps one device copy
-- per host matrix: the key is the 'StableName' of the record's evaluated @values@ vector (the heap object every
-- copy of the record shares), the entry holds the device handle through a 'Weak' pointer keyed on that
-- same vector, so the device copy is finalised (slc_free) when the host matrix is collected.  Every
-- product, transpose and mm below goes through it.
--
-- THREADS.  A handle may be shared by several Haskell threads: the library serialises the calls that
-- build plans or use the handle's transfer pipeline behind a per-handle mutex (csrc/slc_common.cuh,
-- HandleLock), so two sparks forcing @mulV a x@ and @mulV a y@ are safe; they run one after the other.
-- The foreign imports are 'safe' calls for that reason (a blocked call must not stop the capability).
--
-- The functions exported here have the types of their namesakes in Data.Matrix.Sparse
-- (Data/Matrix/Sparse.hs:13-30), so replacing the import is the whole integration:
--
-- > import Data.Matrix.Sparse hiding (mulV, axpy_, axpy, mm, kronecker, transpose, ctrans, hermitian, compress, fromTriples, lin)
-- > import Data.Matrix.Sparse.Device
module Data.Matrix.Sparse.Device
       ( DeviceMatrix, SlcElem
       , toDevice, fromDevice, resident
       , mulV, axpy, axpy_, mulM_, mulMAcc_, mulVD
       , mm, kronecker, transpose, ctrans, hermitian
       , compress, fromTriples, lin
         -- * one process, several GPUs (slc_*_upload_sharded ...)
       , ShardedMatrix, toDevices, mulVS, mulMS_
       ) where

import Control.Concurrent.MVar (MVar, modifyMVar, newMVar)
import Control.Exception (evaluate)
import Control.Monad (when)
import qualified Data.IntMap.Strict as IM
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
import System.Mem.StableName (StableName, eqStableName, hashStableName, makeStableName)
import System.Mem.Weak (Weak, deRefWeak, mkWeak)
import Unsafe.Coerce (unsafeCoerce)

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
type MulM a = Ptr SlcMatrix -> Ptr a -> CLLong -> Ptr a -> CLLong -> CInt -> Ptr a -> IO CInt
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
foreign import ccall "slc_d_kronecker" slc_d_kronecker :: Binary
foreign import ccall "slc_z_kronecker" slc_z_kronecker :: Binary
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
  c_kronecker :: proxy a -> Binary
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
  c_kronecker _ = slc_d_kronecker
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
  c_kronecker _ = slc_z_kronecker
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

-- | The memo for unchanged callers: host matrix -> its device copy (see the module header).
-- Keys are compared with 'eqStableName' inside a bucket of the hash; a dead 'Weak' is dropped when met.
data Entry = forall a. Entry (StableName (Vector a)) (Weak (DeviceMatrix a))

{-# NOINLINE memo #-}
memo :: MVar (IM.IntMap [Entry])
memo = unsafePerformIO (newMVar IM.empty)

-- | The device copy of a host matrix: uploaded on first use, reused afterwards, freed with the matrix.
resident :: forall a. SlcElem a => Matrix Vector a -> IO (DeviceMatrix a)
resident mat = do
  -- The @values@ field is lazy (Data/Matrix/Sparse.hs:68-76), so @values mat@ may still be a selector
  -- thunk: a 'StableName' taken of a thunk differs from the one taken after it has been evaluated, and a
  -- 'Weak' keyed on a thunk dies at the next collection.  Force the field first; key and weak pointer are
  -- then taken of the evaluated vector, the one heap object every later @values mat@ of this record (and of
  -- every copy of it) returns.  Without this the memo may never hit: A is uploaded again per product and
  -- dead entries pile up in their buckets.
  v <- evaluate (values mat)
  key <- makeStableName v
  let h = hashStableName key
  modifyMVar memo $ \tbl -> do
    let bucket = IM.findWithDefault [] h tbl
    alive <- mapM (\e@(Entry k w) -> fmap (\m -> (e, k `eqStableName` key, unsafeCoerce m)) (deRefWeak w)) bucket
    case [d | (_, True, Just d) <- alive] of
      (d : _) -> return (tbl, d :: DeviceMatrix a)
      [] -> do
        d <- toDevice mat
        w <- mkWeak v d Nothing                 -- the ForeignPtr's own finalizer (slc_free) releases the HBM
        let kept = [e | (e, _, Just (_ :: DeviceMatrix a)) <- alive]
        return (IM.insert h (Entry key w : kept) tbl, d)

onDevice :: SlcElem a => Matrix Vector a -> (Ptr SlcMatrix -> IO b) -> IO b
onDevice mat act = do
  DeviceMatrix fp <- resident mat
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
  d <- resident a            -- uploaded once per host matrix, not once per product
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
    c_mulm p px (fromIntegral n) py (fromIntegral n) (fromIntegral k) nullPtr >>= check

-- | Y[:, j] <- A X[:, j] + Y[:, j]: the reference's 'axpy_' into blocks the caller filled
-- (Data/Matrix/Sparse.hs:485-488), beta = 1 of slc_*_mulm.
mulMAcc_ :: SlcElem a => DeviceMatrix a -> Int -> Int -> VSM.IOVector a -> VSM.IOVector a -> IO ()
mulMAcc_ (DeviceMatrix fp) n k xs ys =
  withForeignPtr fp $ \p ->
  VSM.unsafeWith xs $ \px -> VSM.unsafeWith ys $ \py ->
  with 1 $ \pone ->
    c_mulm p px (fromIntegral n) py (fromIntegral n) (fromIntegral k) pone >>= check

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

-- | Kronecker product (kronecker, Data/Matrix/Sparse.hs:597-638): no dimension condition.
kronecker :: forall a. SlcElem a => Matrix Vector a -> Matrix Vector a -> Matrix Vector a
kronecker = binary (c_kronecker ([] :: [a]))

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

-- ---------------------------------------------------------------------------------------------------
-- One process, several GPUs: the sharded handle of include/slc.h.  The matrix is cut into row blocks,
-- one per device, by the library; products take and return whole host vectors.

data SlcSharded
newtype ShardedMatrix a = ShardedMatrix (ForeignPtr SlcSharded)

type UploadS a = CInt -> Ptr CInt -> CInt -> CInt -> Ptr CInt -> Ptr CInt -> Ptr a -> Ptr (Ptr SlcSharded) -> IO CInt
type MulVS a = Ptr SlcSharded -> Ptr a -> Ptr a -> IO CInt
type MulMS a = Ptr SlcSharded -> Ptr a -> CLLong -> Ptr a -> CLLong -> CInt -> IO CInt

foreign import ccall unsafe "&slc_sharded_free" slc_sharded_free :: FunPtr (Ptr SlcSharded -> IO ())
foreign import ccall "slc_d_upload_sharded" slc_d_upload_sharded :: UploadS Double
foreign import ccall "slc_z_upload_sharded" slc_z_upload_sharded :: UploadS (Complex Double)
foreign import ccall "slc_d_mulv_sharded" slc_d_mulv_sharded :: MulVS Double
foreign import ccall "slc_z_mulv_sharded" slc_z_mulv_sharded :: MulVS (Complex Double)
foreign import ccall "slc_d_mulm_sharded" slc_d_mulm_sharded :: MulMS Double
foreign import ccall "slc_z_mulm_sharded" slc_z_mulm_sharded :: MulMS (Complex Double)

class SlcElem a => SlcSharded a where
  c_upload_sharded :: UploadS a
  c_mulv_sharded :: MulVS a
  c_mulm_sharded :: MulMS a

instance SlcSharded Double where
  c_upload_sharded = slc_d_upload_sharded
  c_mulv_sharded = slc_d_mulv_sharded
  c_mulm_sharded = slc_d_mulm_sharded

instance SlcSharded (Complex Double) where
  c_upload_sharded = slc_z_upload_sharded
  c_mulv_sharded = slc_z_mulv_sharded
  c_mulm_sharded = slc_z_mulm_sharded

-- | Cut a host matrix into @ndev@ row blocks on devices 0 .. ndev-1.
toDevices :: SlcSharded a => Int -> Matrix Vector a -> IO (ShardedMatrix a)
toDevices ndev mat =
  withConstMatrix mat $ \nr nc ptrs rows vals ->
  alloca $ \out -> do
    c_upload_sharded (fromIntegral ndev) nullPtr nr nc ptrs rows vals out >>= check
    peek out >>= fmap ShardedMatrix . newForeignPtr slc_sharded_free

-- | A x over all devices (mulV, Data/Matrix/Sparse.hs:464-471).
mulVS :: SlcSharded a => ShardedMatrix a -> Int -> Vector a -> IO (Vector a)
mulVS (ShardedMatrix fp) nr x = withForeignPtr fp $ \p -> do
  ys <- VSM.new nr
  VS.unsafeWith (VS.convert x) $ \px -> VSM.unsafeWith ys $ \py -> c_mulv_sharded p px py >>= check
  VS.convert <$> VS.freeze ys

-- | FEAST's multiplyWork over all devices: k columns of column-major blocks with leading dimension n.
mulMS_ :: SlcSharded a => ShardedMatrix a -> Int -> Int -> VSM.IOVector a -> VSM.IOVector a -> IO ()
mulMS_ (ShardedMatrix fp) n k xs ys =
  withForeignPtr fp $ \p ->
  VSM.unsafeWith xs $ \px -> VSM.unsafeWith ys $ \py ->
    c_mulm_sharded p px (fromIntegral n) py (fromIntegral n) (fromIntegral k) >>= check
