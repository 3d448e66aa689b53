// NativeMethods.cs — P/Invoke binding of liblsmc_b200.so (include/lsmc_b200.h) for the cmdty/storage C# project.
//
// This is the file a maintainer adds under src/Cmdty.Storage/Native/ to call the B200 library from
// LsmcStorageValuation.Calculate (LsmcValuation/LsmcStorageValuation.cs:57); INTEGRATION.md describes the call sequence.
// It cannot be compiled in this repository's image (no .NET toolchain); tests/test_csharp_binding.py parses it and checks
// every struct's field order, field types and resulting size against the ctypes structs the tests run on
// (storage_b200/_abi.py), and that every function the header declares has a DllImport with the same number of parameters.
using System;
using System.Runtime.InteropServices;

namespace Cmdty.Storage.Native
{
    internal enum LsmcStatus
    {
        Ok = 0,
        Argument = 1,               // ArgumentException
        InventoryConstraints = 2,   // InventoryConstraintsCannotBeFulfilledException
        InvalidOperation = 3,       // InvalidOperationException
        Cancelled = 4,              // OperationCanceledException
        Cuda = 5,                   // InvalidOperationException: CUDA failure / no device
        Numerical = 6,              // InvalidOperationException: rank-deficient design matrix
        Unsupported = 7             // NotSupportedException: storage built from opaque Func<> overloads
    }

    internal enum LsmcPanel
    {
        Inventory = 0,
        InjectWithdrawVolume = 1,
        CmdtyConsumed = 2,
        InventoryLoss = 3,
        NetVolume = 4,
        Pv = 5,
        SpotRegression = 6,
        SpotValuation = 7,
        FactorsRegression = 8,
        FactorsValuation = 9,
        DecisionIndex = 10
    }

    internal static unsafe class NativeMethods
    {
        internal const string Lib = "lsmc_b200"; // liblsmc_b200.so

        internal const int MaxBasis = 16, MaxFactors = 8, MaxDecisions = 16, NumTriggerVolumes = 10, CommIdBytes = 128;

        // SimulationDataReturned values are the reference's own (SimulationDataReturned.cs:33-46); the library adds:
        internal const uint ReturnDecisionIndex = 1u << 16, ReferenceOperationOrder = 1u << 17, StreamPanels = 1u << 18;

        [StructLayout(LayoutKind.Sequential)]
        internal struct LsmcStorage
        {
            public int NumPeriods;
            public int MustBeEmptyAtEnd;
            public double* MinInventory;
            public double* MaxInventory;
            public int* ConstraintKind;
            public int* RatchetBegin;
            public int* RatchetCount;
            public int NumRatchetRows;
            public double* RatchetInventory;
            public double* RatchetMinRate;
            public double* RatchetMaxRate;
            public double* InjectionCost;
            public double* InjectionCostDf;
            public double* WithdrawalCost;
            public double* WithdrawalCostDf;
            public double* ConsumedOnInject;
            public double* ConsumedOnWithdraw;
            public double* InventoryLoss;
            public double* InventoryCost;
            public double* InventoryCostDf;
            public fixed double TerminalNpvCoeff[4];
            public double NewtonRaphsonAccuracy;
            public int NewtonRaphsonMaxIterations;
            public int NewtonRaphsonSubdivision;
        }

        [StructLayout(LayoutKind.Sequential)]
        internal struct LsmcProblem
        {
            public LsmcStorage Storage;
            public double Inventory;
            public int CurrentPeriodIsActive;
            public double* ForwardPrice;
            public double* SettleDf;
            public double GridSpacing;
            public double NumericalTolerance;
            public int ExtraDecisions;
            public int DiscountDeltas;
            public uint SimulationDataReturned;
            public int NumFactors;
            public int NumBasis;
            public int* BasisPowers;
            public long NumSims;
            public int ReferenceCompatTerminalPv;
        }

        [StructLayout(LayoutKind.Sequential)]
        internal struct LsmcMultiFactorParams
        {
            public int NumFactors;
            public double* MeanReversion;
            public double* FactorVol;
            public double* FactorCorr;
            public double* PeriodTime;
        }

        [StructLayout(LayoutKind.Sequential)]
        internal struct LsmcTriggerPrices
        {
            public int NumInject;
            public int NumWithdraw;
            public fixed double InjectVolume[10];
            public fixed double InjectPrice[10];
            public fixed double WithdrawVolume[10];
            public fixed double WithdrawPrice[10];
            public int HasInjectPrice;
            public int HasWithdrawPrice;
            public double MaxInjectVolume;
            public double MaxInjectTriggerPrice;
            public double MaxWithdrawVolume;
            public double MaxWithdrawTriggerPrice;
        }

        [StructLayout(LayoutKind.Sequential)]
        internal struct LsmcResults
        {
            public double Npv;
            public double StdError;
            public double BackwardNpv;
            public double* Deltas;
            public double* Profile;
            public LsmcTriggerPrices* TriggerPrices;
            public double* PvBySim;
            public double MsTotal;
            public double MsRegressionSim;
            public double MsValuationSim;
            public double MsBackward;
            public double MsPseudoInverse;
            public double MsForward;
            public long KernelLaunches;
            public double MsBackwardKernel;
            public long BackwardKernelLaunches;
            public double MsDevice;
            public long SnappedInventories;
            public double MsUpload;
        }

        [UnmanagedFunctionPointer(CallingConvention.Cdecl)]
        internal delegate void ProgressFn(double progress, IntPtr user);

        // ---- handle life cycle, inputs
        [DllImport(Lib)] internal static extern int lsmc_create(LsmcProblem* problem, int device, out IntPtr handle);
        [DllImport(Lib)] internal static extern int lsmc_set_sims(IntPtr h, int which, double* spot, double** factors);
        [DllImport(Lib)] internal static extern int lsmc_set_sims_device(IntPtr h, int which, double* spot, double** factors);
        [DllImport(Lib)] internal static extern int lsmc_set_sims_async(IntPtr h, int which, double* spot, double** factors);
        [DllImport(Lib)] internal static extern int lsmc_set_panel_sink(IntPtr h, int which, double* hostDestination);
        [DllImport(Lib)] internal static extern int lsmc_simulate(IntPtr h, int which, LsmcMultiFactorParams* parameters, ulong seed, int antithetic, long firstPathId);
        [DllImport(Lib)] internal static extern int lsmc_simulate_noalloc(int device, LsmcMultiFactorParams* parameters, int numPeriods, int currentPeriodIsActive, double* forwardPrice, long numSims, ulong seed, int antithetic, long firstPathId, double* spotDevice, double** factorsDevice, IntPtr cudaStream);
        [DllImport(Lib)] internal static extern int lsmc_simulate_noalloc_max_periods();
        // ---- the valuation
        [DllImport(Lib)] internal static extern int lsmc_run(IntPtr h, ProgressFn progress, IntPtr user, int* cancel);
        // ---- results
        [DllImport(Lib)] internal static extern int lsmc_num_sim_periods(IntPtr h);
        [DllImport(Lib)] internal static extern int lsmc_num_periods(IntPtr h);
        [DllImport(Lib)] internal static extern int lsmc_get_grid(IntPtr h, int period, double* output, int capacity);
        [DllImport(Lib)] internal static extern long lsmc_num_evaluations(IntPtr h);
        [DllImport(Lib)] internal static extern int lsmc_get_results(IntPtr h, LsmcResults* results);
        [DllImport(Lib)] internal static extern int lsmc_copy_panel(IntPtr h, int which, double* destination);
        [DllImport(Lib)] internal static extern int lsmc_max_grid_points(IntPtr h);
        [DllImport(Lib)] internal static extern int lsmc_get_regression(IntPtr h, double* gamma, double* rInverse);
        [DllImport(Lib)] internal static extern IntPtr lsmc_last_error(IntPtr h);
        [DllImport(Lib)] internal static extern void lsmc_destroy(IntPtr h);
        // ---- paths sharded over several GPUs, one process per GPU
        [DllImport(Lib)] internal static extern int lsmc_comm_unique_id(byte* idOut);
        [DllImport(Lib)] internal static extern int lsmc_comm_init(IntPtr h, byte* id, int rank, int world, long totalSims);
        // ---- host helpers (StorageHelper / grid calc / constraints restated in the library; used by the reference's own unit tests)
        [DllImport(Lib)] internal static extern int lsmc_calculate_inventory_space(LsmcStorage* storage, double startingInventory, double* outMin, double* outMax);
        [DllImport(Lib)] internal static extern int lsmc_bang_bang_decision_set(double minRate, double maxRate, double inventory, double inventoryLoss, double nextMin, double nextMax, double tolerance, int extra, double* output);
        [DllImport(Lib)] internal static extern int lsmc_bisect_inventory_space(double* grid, int n, double inventory, double tolerance, int* lower, int* upper);
        [DllImport(Lib)] internal static extern int lsmc_grid_points(double lower, double upper, double spacing, double* output, int capacity);
        [DllImport(Lib)] internal static extern int lsmc_inject_withdraw_range(LsmcStorage* storage, int period, double inventory, double* minRate, double* maxRate);
        [DllImport(Lib)] internal static extern int lsmc_inventory_space_bounds(LsmcStorage* storage, int period, double nextLower, double nextUpper, double* lower, double* upper);
        [DllImport(Lib, CharSet = CharSet.Ansi)] internal static extern int lsmc_parse_basis(string expression, int numFactors, int* powers, int capacity);

        /// <summary>Maps a status to the exception the reference throws on the same condition (include/lsmc_b200.h lsmc_status).</summary>
        internal static void Check(int status, IntPtr handle)
        {
            if (status == (int)LsmcStatus.Ok) return;
            string message = Marshal.PtrToStringAnsi(lsmc_last_error(handle)) ?? ("lsmc status " + status);
            switch ((LsmcStatus)status)
            {
                case LsmcStatus.Argument: throw new ArgumentException(message);
                case LsmcStatus.InventoryConstraints: throw new InventoryConstraintsCannotBeFulfilledException(message);
                case LsmcStatus.Cancelled: throw new OperationCanceledException(message);
                case LsmcStatus.Unsupported: throw new NotSupportedException(message);
                default: throw new InvalidOperationException(message);
            }
        }
    }
}
